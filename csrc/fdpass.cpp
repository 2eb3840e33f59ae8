// File-descriptor passing between the ranks of one host (SCM_RIGHTS over abstract Unix sockets).
// CUDA virtual-memory allocations and NVSwitch multicast objects are shared between processes as
// POSIX file descriptors; the rendezvous store can carry bytes but not descriptors, so each rank
// listens on "\0b200dp-<token>-<rank>" and peers connect, send one descriptor plus an integer tag,
// and close.  (No counterpart in the reference: its gloo backend moves bytes over TCP only.)
#include <errno.h>
#include <poll.h>
#include <stdio.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <chrono>
#include <string>
#include <thread>

#include "host_util.h"

namespace b200dp {

static socklen_t make_addr(sockaddr_un* addr, const std::string& name) {
  memset(addr, 0, sizeof(*addr));
  addr->sun_family = AF_UNIX;
  // abstract namespace: leading NUL, no filesystem entry to clean up
  const size_t n = name.size() < sizeof(addr->sun_path) - 2 ? name.size() : sizeof(addr->sun_path) - 2;
  memcpy(addr->sun_path + 1, name.data(), n);
  return static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + n);
}

int fd_listen(const std::string& name) {
  int s = ::socket(AF_UNIX, SOCK_STREAM, 0);
  if (s < 0) return -1;
  sockaddr_un addr;
  const socklen_t len = make_addr(&addr, name);
  if (::bind(s, reinterpret_cast<sockaddr*>(&addr), len) != 0 || ::listen(s, 32) != 0) {
    ::close(s);
    return -1;
  }
  return s;
}

// Connect to `name` (retrying for up to timeout_ms), send `fd` with `tag` as payload.
int fd_send(const std::string& name, int fd, int tag, int timeout_ms) {
  sockaddr_un addr;
  const socklen_t len = make_addr(&addr, name);
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  int s = -1;
  for (;;) {
    s = ::socket(AF_UNIX, SOCK_STREAM, 0);
    if (s >= 0 && ::connect(s, reinterpret_cast<sockaddr*>(&addr), len) == 0) break;
    if (s >= 0) ::close(s);
    s = -1;
    if (std::chrono::steady_clock::now() > deadline) return -1;
    std::this_thread::sleep_for(std::chrono::milliseconds(20));
  }
  msghdr msg{};
  iovec iov{&tag, sizeof(tag)};
  char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  const ssize_t k = ::sendmsg(s, &msg, 0);
  ::close(s);
  return k == static_cast<ssize_t>(sizeof(tag)) ? 0 : -1;
}

// Accept one connection on `listen_fd` and return the descriptor it carries (tag through *tag).
// timeout_ms < 0 waits for ever; otherwise -2 is returned when nobody connected in time.
int fd_recv_timeout(int listen_fd, int* tag, int timeout_ms) {
  if (timeout_ms >= 0) {
    pollfd pfd{listen_fd, POLLIN, 0};
    int rc;
    do {
      rc = ::poll(&pfd, 1, timeout_ms);
    } while (rc < 0 && errno == EINTR);
    if (rc == 0) return -2;
    if (rc < 0) return -1;
  }
  int s = ::accept(listen_fd, nullptr, nullptr);
  if (s < 0) return -1;
  msghdr msg{};
  int payload = -1;
  iovec iov{&payload, sizeof(payload)};
  char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  const ssize_t k = ::recvmsg(s, &msg, 0);
  ::close(s);
  if (k != static_cast<ssize_t>(sizeof(payload))) return -1;
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (c == nullptr || c->cmsg_level != SOL_SOCKET || c->cmsg_type != SCM_RIGHTS) return -1;
  int fd = -1;
  memcpy(&fd, CMSG_DATA(c), sizeof(int));
  if (tag) *tag = payload;
  return fd;
}

int fd_recv(int listen_fd, int* tag) { return fd_recv_timeout(listen_fd, tag, -1); }

}  // namespace b200dp

// C entry points: the descriptor hand-off is host-only code, so it can be exercised without a GPU
// (tests/test_cpu_runtime.py).  name: socket name in the abstract namespace, without the leading NUL.
extern "C" int b200dp_fd_listen(const char* name) { return name ? b200dp::fd_listen(name) : -1; }
extern "C" int b200dp_fd_send(const char* name, int fd, int tag, int timeout_ms) {
  return name ? b200dp::fd_send(name, fd, tag, timeout_ms) : -1;
}
/* Returns the received descriptor (>= 0), -2 on timeout, -1 on error.  timeout_ms < 0: wait for ever. */
extern "C" int b200dp_fd_recv(int listen_fd, int* tag, int timeout_ms) {
  return b200dp::fd_recv_timeout(listen_fd, tag, timeout_ms);
}
