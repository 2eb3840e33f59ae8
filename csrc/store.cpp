// Rendezvous key-value store over TCP: rank 0 hosts it on (master_ip, port), every rank
// (rank 0 included) connects as a client.  It carries the few hundred bytes the ranks must
// exchange before they can talk over NVLink: CUDA IPC memory handles of the symmetric heaps,
// barrier counters, and small control values.
//
// This is the framework-owned stand-in for what the reference obtains from c10d's TCPStore
// through dist.init_process_group with MASTER_ADDR / MASTER_PORT
// (ref: master/part2a/part2a.py:80-85).
//
// Wire format (all integers little-endian):
//   request : u8 op | u32 key_len | key bytes | u32 val_len | val bytes
//   reply   : u32 status (0 ok, 1 timeout) | u32 val_len | val bytes
// ops: 1 SET, 2 GET (blocks until the key exists or timeout), 3 ADD (i64 delta -> new value),
//      4 DEL (erase every key with the given prefix)
#include <arpa/inet.h>
#include <errno.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <stdint.h>
#include <string.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../include/b200dp.h"
#include "host_util.h"

namespace b200dp {

enum : uint8_t { OP_SET = 1, OP_GET = 2, OP_ADD = 3, OP_DEL = 4 };

static bool send_all(int fd, const void* buf, size_t n) {
  const char* p = static_cast<const char*>(buf);
  while (n > 0) {
    ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
    if (k <= 0) {
      if (k < 0 && errno == EINTR) continue;
      return false;
    }
    p += k;
    n -= static_cast<size_t>(k);
  }
  return true;
}

static bool recv_all(int fd, void* buf, size_t n) {
  char* p = static_cast<char*>(buf);
  while (n > 0) {
    ssize_t k = ::recv(fd, p, n, 0);
    if (k <= 0) {
      if (k < 0 && errno == EINTR) continue;
      return false;
    }
    p += k;
    n -= static_cast<size_t>(k);
  }
  return true;
}

class StoreServer {
 public:
  StoreServer() = default;
  ~StoreServer() { stop(); }

  // Binds to `ip` when it is a numeric address of this host (so the store is not reachable through other
  // interfaces); a host name or an address that cannot be bound (NAT, 0.0.0.0) falls back to every interface.
  bool start(const char* ip, int port, std::string* err) {
    listen_fd_ = ::socket(AF_INET, SOCK_STREAM, 0);
    if (listen_fd_ < 0) {
      *err = "socket() failed";
      return false;
    }
    int one = 1;
    setsockopt(listen_fd_, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    addr.sin_port = htons(static_cast<uint16_t>(port));
    in_addr want{};
    bool bound = false;
    if (ip != nullptr && inet_pton(AF_INET, ip, &want) == 1 && want.s_addr != htonl(INADDR_ANY)) {
      addr.sin_addr = want;
      bound = ::bind(listen_fd_, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) == 0;
    }
    addr.sin_addr.s_addr = htonl(INADDR_ANY);
    if (!bound && ::bind(listen_fd_, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0) {
      *err = std::string("bind() failed: ") + strerror(errno);
      ::close(listen_fd_);
      listen_fd_ = -1;
      return false;
    }
    if (::listen(listen_fd_, 64) != 0) {
      *err = "listen() failed";
      ::close(listen_fd_);
      listen_fd_ = -1;
      return false;
    }
    accept_thread_ = std::thread([this] { accept_loop(); });
    return true;
  }

  void stop() {
    if (stopping_.exchange(true)) return;
    if (listen_fd_ >= 0) {
      ::shutdown(listen_fd_, SHUT_RDWR);
      ::close(listen_fd_);
    }
    {
      std::lock_guard<std::mutex> lk(mu_);
      for (int fd : client_fds_) ::shutdown(fd, SHUT_RDWR);
    }
    cv_.notify_all();
    if (accept_thread_.joinable()) accept_thread_.join();
    for (auto& t : workers_)
      if (t.joinable()) t.join();
  }

 private:
  void accept_loop() {
    while (!stopping_) {
      int fd = ::accept(listen_fd_, nullptr, nullptr);
      if (fd < 0) {
        if (stopping_) break;
        continue;
      }
      int one = 1;
      setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
      std::lock_guard<std::mutex> lk(mu_);
      client_fds_.push_back(fd);
      workers_.emplace_back([this, fd] { serve(fd); });
    }
  }

  // A request that claims a key above 4 KiB or a value above 1 MiB is not one of ours: drop the connection
  // instead of allocating what it asks for (the port is reachable by anything that can reach the host).
  static constexpr uint32_t kMaxKeyLen = 4096, kMaxValLen = 1u << 20;

  void serve(int fd) {
    try {
      serve_loop(fd);
    } catch (...) {
      // e.g. std::bad_alloc: this client is dropped, the store keeps serving the others
    }
    ::close(fd);
  }

  void serve_loop(int fd) {
    for (;;) {
      uint8_t op;
      uint32_t klen, vlen;
      if (!recv_all(fd, &op, 1) || !recv_all(fd, &klen, 4)) break;
      if (klen > kMaxKeyLen) break;
      std::string key(klen, '\0');
      if (klen && !recv_all(fd, &key[0], klen)) break;
      if (!recv_all(fd, &vlen, 4)) break;
      if (vlen > kMaxValLen) break;
      std::string val(vlen, '\0');
      if (vlen && !recv_all(fd, &val[0], vlen)) break;
      uint32_t status = 0;
      std::string out;
      if (op == OP_SET) {
        std::lock_guard<std::mutex> lk(mu_);
        kv_[key] = val;
        cv_.notify_all();
      } else if (op == OP_GET) {
        int64_t timeout_ms = 0;
        if (vlen == 8) memcpy(&timeout_ms, val.data(), 8);
        std::unique_lock<std::mutex> lk(mu_);
        auto pred = [&] { return stopping_.load() || kv_.count(key) > 0; };
        if (timeout_ms > 0) {
          if (!cv_.wait_for(lk, std::chrono::milliseconds(timeout_ms), pred)) status = 1;
        } else {
          cv_.wait(lk, pred);
        }
        if (status == 0 && kv_.count(key)) out = kv_[key];
        else status = 1;
      } else if (op == OP_ADD) {
        int64_t delta = 0;
        if (vlen == 8) memcpy(&delta, val.data(), 8);
        std::lock_guard<std::mutex> lk(mu_);
        int64_t cur = 0;
        auto it = kv_.find(key);
        if (it != kv_.end() && it->second.size() == 8) memcpy(&cur, it->second.data(), 8);
        cur += delta;
        std::string enc(8, '\0');
        memcpy(&enc[0], &cur, 8);
        kv_[key] = enc;
        out = enc;
        cv_.notify_all();
      } else if (op == OP_DEL) {
        // erase every key that starts with `key` (finished barrier rounds, host-side reductions)
        std::lock_guard<std::mutex> lk(mu_);
        for (auto it = kv_.lower_bound(key); it != kv_.end() && it->first.compare(0, key.size(), key) == 0;)
          it = kv_.erase(it);
      } else {
        break;
      }
      uint32_t olen = static_cast<uint32_t>(out.size());
      if (!send_all(fd, &status, 4) || !send_all(fd, &olen, 4) || (olen && !send_all(fd, out.data(), olen))) break;
    }
  }

  int listen_fd_ = -1;
  std::atomic<bool> stopping_{false};
  std::thread accept_thread_;
  std::vector<std::thread> workers_;
  std::vector<int> client_fds_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::map<std::string, std::string> kv_;
};

struct Store {
  int rank = 0, world = 1;
  int fd = -1;
  int timeout_ms = 120000;
  std::mutex mu;  // one request in flight per client
  StoreServer* server = nullptr;
  int64_t barrier_round = 0;

  bool request(uint8_t op, const std::string& key, const void* val, uint32_t vlen, uint32_t* status,
               std::string* out) {
    std::lock_guard<std::mutex> lk(mu);
    uint32_t klen = static_cast<uint32_t>(key.size());
    if (!send_all(fd, &op, 1) || !send_all(fd, &klen, 4) || !send_all(fd, key.data(), klen) ||
        !send_all(fd, &vlen, 4) || (vlen && !send_all(fd, val, vlen)))
      return false;
    uint32_t olen = 0;
    if (!recv_all(fd, status, 4) || !recv_all(fd, &olen, 4)) return false;
    out->resize(olen);
    if (olen && !recv_all(fd, &(*out)[0], olen)) return false;
    return true;
  }
};

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_store_create(const char* master_ip, int port, int rank, int world, int timeout_ms,
                                   void** out_handle) {
  if (!master_ip || !out_handle || world < 1 || rank < 0 || rank >= world) {
    set_error("store_create: bad arguments (rank %d of %d)", rank, world);
    return B200DP_EINVAL;
  }
  auto* s = new Store();
  s->rank = rank;
  s->world = world;
  s->timeout_ms = timeout_ms > 0 ? timeout_ms : 120000;
  if (rank == 0) {
    s->server = new StoreServer();
    std::string err;
    if (!s->server->start(master_ip, port, &err)) {
      set_error("store_create: rank 0 cannot host the store on port %d: %s", port, err.c_str());
      delete s->server;
      delete s;
      return B200DP_EINVAL;
    }
  }
  addrinfo hints{}, *res = nullptr;
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  const std::string port_s = std::to_string(port);
  if (getaddrinfo(master_ip, port_s.c_str(), &hints, &res) != 0 || res == nullptr) {
    set_error("store_create: cannot resolve master address %s", master_ip);
    if (s->server) delete s->server;
    delete s;
    return B200DP_EINVAL;
  }
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(s->timeout_ms);
  int fd = -1;
  while (true) {
    fd = ::socket(AF_INET, SOCK_STREAM, 0);
    if (fd >= 0 && ::connect(fd, res->ai_addr, res->ai_addrlen) == 0) break;
    if (fd >= 0) ::close(fd);
    fd = -1;
    if (std::chrono::steady_clock::now() > deadline) break;
    std::this_thread::sleep_for(std::chrono::milliseconds(50));
  }
  freeaddrinfo(res);
  if (fd < 0) {
    set_error("store_create: rank %d could not reach the store at %s:%d within %d ms", rank, master_ip, port,
              s->timeout_ms);
    if (s->server) delete s->server;
    delete s;
    return B200DP_ETIMEDOUT;
  }
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
  s->fd = fd;
  *out_handle = s;
  return 0;
}

extern "C" int b200dp_store_set(void* h, const char* key, const void* data, int len) {
  auto* s = static_cast<Store*>(h);
  uint32_t status;
  std::string out;
  if (!s->request(OP_SET, key, data, static_cast<uint32_t>(len), &status, &out)) {
    set_error("store_set(%s): connection to the store lost", key);
    return B200DP_ETIMEDOUT;
  }
  return 0;
}

extern "C" int b200dp_store_get(void* h, const char* key, void* buf, int cap, int* out_len) {
  auto* s = static_cast<Store*>(h);
  uint32_t status;
  std::string out;
  int64_t t = s->timeout_ms;
  if (!s->request(OP_GET, key, &t, 8, &status, &out)) {
    set_error("store_get(%s): connection to the store lost", key);
    return B200DP_ETIMEDOUT;
  }
  if (status != 0) {
    set_error("store_get(%s): rank %d timed out after %d ms waiting for the key (is a peer rank down?)", key,
              s->rank, s->timeout_ms);
    return B200DP_ETIMEDOUT;
  }
  if (static_cast<int>(out.size()) > cap) {
    set_error("store_get(%s): value of %zu bytes exceeds the %d-byte buffer", key, out.size(), cap);
    return B200DP_EINVAL;
  }
  memcpy(buf, out.data(), out.size());
  if (out_len) *out_len = static_cast<int>(out.size());
  return 0;
}

extern "C" int b200dp_store_add(void* h, const char* key, long long delta, long long* result) {
  auto* s = static_cast<Store*>(h);
  uint32_t status;
  std::string out;
  int64_t d = delta;
  if (!s->request(OP_ADD, key, &d, 8, &status, &out) || out.size() != 8) {
    set_error("store_add(%s): connection to the store lost", key);
    return B200DP_ETIMEDOUT;
  }
  int64_t v;
  memcpy(&v, out.data(), 8);
  if (result) *result = v;
  return 0;
}

extern "C" int b200dp_store_delete_prefix(void* h, const char* prefix) {
  auto* s = static_cast<Store*>(h);
  uint32_t status;
  std::string out;
  if (!prefix || !prefix[0] || !s->request(OP_DEL, prefix, nullptr, 0, &status, &out)) {
    set_error("store_delete_prefix(%s): connection to the store lost", prefix ? prefix : "(null)");
    return B200DP_ETIMEDOUT;
  }
  return 0;
}

extern "C" int b200dp_store_timeout_ms(void* h) {
  auto* s = static_cast<Store*>(h);
  return s ? s->timeout_ms : 0;
}

// All ranks call this the same number of times (the names need not match: rounds are numbered).
extern "C" int b200dp_store_barrier(void* h, const char* name) {
  auto* s = static_cast<Store*>(h);
  const int64_t round = s->barrier_round++;
  (void)name;  // the round number alone identifies the barrier, so ranks on different code paths cannot pair up wrongly
  const std::string base = std::string("barrier/") + std::to_string(round);
  // everybody passed round - 2 before anyone could arrive here, so its keys are dead: rank 0 sweeps them
  if (s->rank == 0 && round >= 2) b200dp_store_delete_prefix(h, ("barrier/" + std::to_string(round - 2) + "/").c_str());
  long long arrived = 0;
  int rc = b200dp_store_add(h, (base + "/count").c_str(), 1, &arrived);
  if (rc) return rc;
  if (arrived == s->world) {
    const char one = 1;
    rc = b200dp_store_set(h, (base + "/go").c_str(), &one, 1);
    if (rc) return rc;
  }
  char tmp[8];
  int len = 0;
  return b200dp_store_get(h, (base + "/go").c_str(), tmp, sizeof(tmp), &len);
}

extern "C" int b200dp_store_destroy(void* h) {
  auto* s = static_cast<Store*>(h);
  if (!s) return 0;
  if (s->fd >= 0) {
    ::shutdown(s->fd, SHUT_RDWR);
    ::close(s->fd);
  }
  if (s->server) {
    s->server->stop();
    delete s->server;
  }
  delete s;
  return 0;
}
