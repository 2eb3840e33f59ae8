// Error buffer + version string of libb200dp.so.
#include "../include/b200dp.h"
#include <stdlib.h>

#include "host_util.h"

namespace b200dp {
char* error_buffer() {
  static thread_local char buf[kErrorBufferSize] = {0};
  return buf;
}
bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("B200DP_PDL");
    return e == nullptr || e[0] != '0';
  }();
  return on;
}
}  // namespace b200dp

extern "C" const char* b200dp_last_error(void) { return b200dp::error_buffer(); }
extern "C" const char* b200dp_version(void) { return "b200dp 0.1 (sm_100a; tcgen05/TMEM/TMA; NVLink P2P collectives)"; }
