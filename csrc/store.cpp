// R1 — bootstrap key-value store.
//
// A small TCP KV service hosted by rank 0 on MASTER_ADDR:<port>. It is the only
// control-plane transport of the native backend: it carries the rendezvous of a
// comm (unix-socket names, multicast readiness) and host barriers. The data
// plane never touches it.
//
// Role parity: torch TCPStore/PrefixStore as reached from the reference at
// main.py:92-94 ([torch] distributed/rendezvous.py:162-208). Prefixing is done by
// the callers (every comm key starts with its group key).
//
// Wire format (little endian, one request -> one reply, per connection in order):
//   request : u8 op | u32 klen | key | u32 vlen | val | i64 arg
//   reply   : i32 status | u32 vlen | val | i64 num
// The server runs one thread per connection; blocking GETs wait on a condition
// variable, so a waiting client never holds up another one.

#include <arpa/inet.h>
#include <errno.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
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

#include "common.h"

namespace {

enum : uint8_t { OP_SET = 1, OP_GET = 2, OP_ADD = 3, OP_CHECK = 4, OP_DEL = 5 };

bool send_all(int fd, const void* p, size_t n) {
  const char* c = static_cast<const char*>(p);
  while (n) {
    ssize_t k = ::send(fd, c, n, MSG_NOSIGNAL);
    if (k < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    c += k;
    n -= static_cast<size_t>(k);
  }
  return true;
}

bool recv_all(int fd, void* p, size_t n) {
  char* c = static_cast<char*>(p);
  while (n) {
    ssize_t k = ::recv(fd, c, n, 0);
    if (k == 0) return false;
    if (k < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    c += k;
    n -= static_cast<size_t>(k);
  }
  return true;
}

}  // namespace

struct pdcc_store_server {
  int listen_fd = -1;
  int port = 0;
  std::atomic<bool> stopping{false};
  std::thread accept_thread;
  std::mutex mu;
  std::condition_variable cv;
  std::map<std::string, std::string> kv;
  std::vector<std::thread> workers;
  std::vector<int> client_fds;

  void serve(int fd) {
    for (;;) {
      uint8_t op;
      uint32_t klen, vlen;
      int64_t arg;
      if (!recv_all(fd, &op, 1) || !recv_all(fd, &klen, 4)) break;
      if (klen > (1u << 20)) break;
      std::string key(klen, '\0');
      if (klen && !recv_all(fd, &key[0], klen)) break;
      if (!recv_all(fd, &vlen, 4)) break;
      if (vlen > (64u << 20)) break;
      std::string val(vlen, '\0');
      if (vlen && !recv_all(fd, &val[0], vlen)) break;
      if (!recv_all(fd, &arg, 8)) break;

      int32_t status = 0;
      std::string out;
      int64_t num = 0;
      switch (op) {
        case OP_SET: {
          std::lock_guard<std::mutex> g(mu);
          kv[key] = std::move(val);
          cv.notify_all();
          break;
        }
        case OP_GET: {  // arg = timeout in ms (<0: wait forever)
          std::unique_lock<std::mutex> g(mu);
          auto ready = [&] { return stopping.load() || kv.count(key) > 0; };
          if (arg < 0) {
            cv.wait(g, ready);
          } else {
            cv.wait_for(g, std::chrono::milliseconds(arg), ready);
          }
          auto it = kv.find(key);
          if (it == kv.end()) {
            status = PDCC_ETIMEOUT;
          } else {
            out = it->second;
          }
          break;
        }
        case OP_ADD: {
          std::lock_guard<std::mutex> g(mu);
          int64_t cur = 0;
          auto it = kv.find(key);
          if (it != kv.end() && it->second.size() == 8) memcpy(&cur, it->second.data(), 8);
          cur += arg;
          kv[key] = std::string(reinterpret_cast<const char*>(&cur), 8);
          num = cur;
          cv.notify_all();
          break;
        }
        case OP_CHECK: {
          std::lock_guard<std::mutex> g(mu);
          num = kv.count(key) ? 1 : 0;
          break;
        }
        case OP_DEL: {
          std::lock_guard<std::mutex> g(mu);
          num = static_cast<int64_t>(kv.erase(key));
          break;
        }
        default:
          status = PDCC_EINVAL;
      }
      uint32_t olen = static_cast<uint32_t>(out.size());
      if (!send_all(fd, &status, 4) || !send_all(fd, &olen, 4) ||
          (olen && !send_all(fd, out.data(), olen)) || !send_all(fd, &num, 8))
        break;
    }
    {
      // forget the descriptor before closing it: server_stop must never shutdown() a number
      // the process has meanwhile reused for something else
      std::lock_guard<std::mutex> g(mu);
      for (auto& c : client_fds)
        if (c == fd) c = -1;
    }
    ::close(fd);
  }

  void accept_loop() {
    while (!stopping.load()) {
      pollfd p{listen_fd, POLLIN, 0};
      int r = ::poll(&p, 1, 200);
      if (r <= 0) continue;
      int fd = ::accept(listen_fd, nullptr, nullptr);
      if (fd < 0) continue;
      int one = 1;
      setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
      std::lock_guard<std::mutex> g(mu);
      client_fds.push_back(fd);
      workers.emplace_back([this, fd] { serve(fd); });
    }
  }
};

struct pdcc_store {
  int fd = -1;
  std::mutex mu;  // one request in flight per client

  int request(uint8_t op, const char* key, const void* val, uint32_t vlen, int64_t arg,
              std::string* out, int64_t* num) {
    std::lock_guard<std::mutex> g(mu);
    uint32_t klen = static_cast<uint32_t>(strlen(key));
    std::string msg;
    msg.reserve(1 + 4 + klen + 4 + vlen + 8);
    msg.append(reinterpret_cast<const char*>(&op), 1);
    msg.append(reinterpret_cast<const char*>(&klen), 4);
    msg.append(key, klen);
    msg.append(reinterpret_cast<const char*>(&vlen), 4);
    if (vlen) msg.append(static_cast<const char*>(val), vlen);
    msg.append(reinterpret_cast<const char*>(&arg), 8);
    if (!send_all(fd, msg.data(), msg.size())) {
      pdcc::set_error("store: send failed: %s", strerror(errno));
      return PDCC_ESYS;
    }
    int32_t status;
    uint32_t olen;
    int64_t n;
    if (!recv_all(fd, &status, 4) || !recv_all(fd, &olen, 4)) {
      pdcc::set_error("store: connection closed by server");
      return PDCC_ESYS;
    }
    std::string o(olen, '\0');
    if (olen && !recv_all(fd, &o[0], olen)) {
      pdcc::set_error("store: short reply");
      return PDCC_ESYS;
    }
    if (!recv_all(fd, &n, 8)) {
      pdcc::set_error("store: short reply");
      return PDCC_ESYS;
    }
    if (out) *out = std::move(o);
    if (num) *num = n;
    if (status == PDCC_ETIMEOUT) pdcc::set_error("store: timed out waiting for key '%s'", key);
    return status;
  }
};

extern "C" {

int pdcc_store_server_start(const char* bind_addr, int port, pdcc_store_server_t** out,
                            int* bound_port) {
  if (!out) return PDCC_EINVAL;
  int fd = ::socket(AF_INET, SOCK_STREAM, 0);
  if (fd < 0) {
    pdcc::set_error("store: socket(): %s", strerror(errno));
    return PDCC_ESYS;
  }
  int one = 1;
  setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_port = htons(static_cast<uint16_t>(port));
  if (!bind_addr || !*bind_addr || inet_pton(AF_INET, bind_addr, &addr.sin_addr) != 1)
    addr.sin_addr.s_addr = htonl(INADDR_ANY);
  if (::bind(fd, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0) {
    pdcc::set_error("store: bind(%d): %s", port, strerror(errno));
    ::close(fd);
    return PDCC_ESYS;
  }
  if (::listen(fd, 128) != 0) {
    pdcc::set_error("store: listen(): %s", strerror(errno));
    ::close(fd);
    return PDCC_ESYS;
  }
  socklen_t alen = sizeof(addr);
  getsockname(fd, reinterpret_cast<sockaddr*>(&addr), &alen);
  auto* s = new pdcc_store_server();
  s->listen_fd = fd;
  s->port = ntohs(addr.sin_port);
  s->accept_thread = std::thread([s] { s->accept_loop(); });
  if (bound_port) *bound_port = s->port;
  *out = s;
  return PDCC_OK;
}

int pdcc_store_server_stop(pdcc_store_server_t* s) {
  if (!s) return PDCC_EINVAL;
  s->stopping.store(true);
  if (s->accept_thread.joinable()) s->accept_thread.join();
  {
    std::lock_guard<std::mutex> g(s->mu);
    for (int fd : s->client_fds)
      if (fd >= 0) ::shutdown(fd, SHUT_RDWR);
    s->cv.notify_all();
  }
  for (auto& t : s->workers)
    if (t.joinable()) t.join();
  ::close(s->listen_fd);
  delete s;
  return PDCC_OK;
}

int pdcc_store_connect(const char* host, int port, int timeout_ms, pdcc_store_t** out) {
  if (!out || !host) return PDCC_EINVAL;
  addrinfo hints{};
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  addrinfo* res = nullptr;
  char portstr[16];
  snprintf(portstr, sizeof(portstr), "%d", port);
  if (getaddrinfo(host, portstr, &hints, &res) != 0 || !res) {
    pdcc::set_error("store: cannot resolve %s", host);
    return PDCC_ESYS;
  }
  auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  int fd = -1;
  // The server may not be up yet (rank 0 is still importing torch): retry.
  for (;;) {
    fd = ::socket(AF_INET, SOCK_STREAM, 0);
    if (fd >= 0 && ::connect(fd, res->ai_addr, res->ai_addrlen) == 0) break;
    if (fd >= 0) ::close(fd);
    fd = -1;
    if (std::chrono::steady_clock::now() > deadline) break;
    std::this_thread::sleep_for(std::chrono::milliseconds(20));
  }
  freeaddrinfo(res);
  if (fd < 0) {
    pdcc::set_error("store: connect(%s:%d) timed out after %d ms", host, port, timeout_ms);
    return PDCC_ETIMEOUT;
  }
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
  auto* st = new pdcc_store();
  st->fd = fd;
  *out = st;
  return PDCC_OK;
}

int pdcc_store_close(pdcc_store_t* st) {
  if (!st) return PDCC_EINVAL;
  ::close(st->fd);
  delete st;
  return PDCC_OK;
}

int pdcc_store_set(pdcc_store_t* st, const char* key, const void* val, int len) {
  if (!st || !key || len < 0) return PDCC_EINVAL;
  return st->request(OP_SET, key, val, static_cast<uint32_t>(len), 0, nullptr, nullptr);
}

int pdcc_store_get(pdcc_store_t* st, const char* key, void* buf, int cap, int* len_out,
                   int timeout_ms) {
  if (!st || !key) return PDCC_EINVAL;
  std::string out;
  int rc = st->request(OP_GET, key, nullptr, 0, timeout_ms, &out, nullptr);
  if (rc != PDCC_OK) return rc;
  if (len_out) *len_out = static_cast<int>(out.size());
  if (buf && cap > 0) memcpy(buf, out.data(), std::min<size_t>(out.size(), cap));
  return PDCC_OK;
}

int pdcc_store_add(pdcc_store_t* st, const char* key, long long delta, long long* result) {
  if (!st || !key) return PDCC_EINVAL;
  int64_t n = 0;
  int rc = st->request(OP_ADD, key, nullptr, 0, delta, nullptr, &n);
  if (result) *result = n;
  return rc;
}

int pdcc_store_check(pdcc_store_t* st, const char* key, int* exists) {
  if (!st || !key) return PDCC_EINVAL;
  int64_t n = 0;
  int rc = st->request(OP_CHECK, key, nullptr, 0, 0, nullptr, &n);
  if (exists) *exists = static_cast<int>(n);
  return rc;
}

int pdcc_store_delete(pdcc_store_t* st, const char* key) {
  if (!st || !key) return PDCC_EINVAL;
  return st->request(OP_DEL, key, nullptr, 0, 0, nullptr, nullptr);
}

int pdcc_store_barrier(pdcc_store_t* st, const char* name, int nranks, int timeout_ms) {
  if (!st || !name || nranks < 1) return PDCC_EINVAL;
  std::string cnt = std::string(name) + "/count";
  std::string done = std::string(name) + "/done";
  long long n = 0;
  int rc = pdcc_store_add(st, cnt.c_str(), 1, &n);
  if (rc != PDCC_OK) return rc;
  if (n == nranks) {
    rc = pdcc_store_set(st, done.c_str(), "1", 1);
    if (rc != PDCC_OK) return rc;
  }
  return pdcc_store_get(st, done.c_str(), nullptr, 0, nullptr, timeout_ms);
}

}  // extern "C"
