// TCP key-value store: the rendezvous backbone of b200dp (W1 in SURVEY.md section 8).
//
// The reference gets this from torch: `dist.init_process_group(backend, rank=, world_size=)`
// with the implicit env:// method builds a C++ TCPStore on MASTER_ADDR:MASTER_PORT hosted by
// rank 0 (test_init.py:76-91, allreduce_toy.py:52-58, mnist_distributed.py:124-125 and 50).
// This is our own store: one poll()-driven server thread (rank 0) and blocking clients.
// Beyond set/get it offers the primitives the process-group layer needs: atomic add (rank
// counting, barriers), blocking wait-with-timeout (a late or dead peer becomes an error
// message instead of a hang), compare-free delete and key counting.
//
// Wire format (little endian): request  = u8 op | u32 klen | key | u32 vlen | value | i64 arg
//                              response = i32 status | u32 vlen | value | i64 num
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#define B200H_API extern "C" __attribute__((visibility("default")))

namespace {

thread_local char g_err[1024];
int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

enum Op : uint8_t { OP_SET = 1, OP_GET = 2, OP_ADD = 3, OP_CHECK = 4, OP_DELETE = 5, OP_NUMKEYS = 6, OP_PING = 7 };
enum Status : int32_t { ST_OK = 0, ST_TIMEOUT = 1, ST_MISSING = 2, ST_BAD = 3 };

using Clock = std::chrono::steady_clock;

bool send_all(int fd, const void* buf, size_t n) {
  const char* p = static_cast<const char*>(buf);
  while (n > 0) {
    ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
    if (k < 0) { if (errno == EINTR) continue; return false; }
    p += k; n -= (size_t)k;
  }
  return true;
}
bool recv_all(int fd, void* buf, size_t n) {
  char* p = static_cast<char*>(buf);
  while (n > 0) {
    ssize_t k = ::recv(fd, p, n, 0);
    if (k == 0) return false;
    if (k < 0) { if (errno == EINTR) continue; return false; }
    p += k; n -= (size_t)k;
  }
  return true;
}

struct Request {
  uint8_t op = 0;
  std::string key, val;
  int64_t arg = 0;
};

struct Conn {
  int fd = -1;
  std::vector<char> inbuf;
};

struct Waiter {
  int fd;
  std::string key;
  Clock::time_point deadline;
  bool has_deadline;
};

class Server {
 public:
  int start(const char* host, int port) {
    listen_fd_ = ::socket(AF_INET, SOCK_STREAM, 0);
    if (listen_fd_ < 0) return fail(-1, "store: socket() failed: %s", strerror(errno));
    int one = 1;
    // SO_REUSEADDR BEFORE bind (the reference's find_free_port sets it after, test_init.py:51-52)
    setsockopt(listen_fd_, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    addr.sin_port = htons((uint16_t)port);
    if (!host || !*host || inet_pton(AF_INET, host, &addr.sin_addr) != 1) addr.sin_addr.s_addr = htonl(INADDR_ANY);
    if (::bind(listen_fd_, (sockaddr*)&addr, sizeof(addr)) != 0) {
      int e = errno; ::close(listen_fd_); listen_fd_ = -1;
      return fail(-1, "store: cannot bind %s:%d: %s", host ? host : "*", port, strerror(e));
    }
    if (::listen(listen_fd_, 512) != 0) {
      int e = errno; ::close(listen_fd_); listen_fd_ = -1;
      return fail(-1, "store: listen failed: %s", strerror(e));
    }
    socklen_t len = sizeof(addr);
    getsockname(listen_fd_, (sockaddr*)&addr, &len);
    port_ = ntohs(addr.sin_port);
    if (::pipe(wake_) != 0) return fail(-1, "store: pipe failed");
    running_ = true;
    thread_ = std::thread([this] { loop(); });
    return 0;
  }
  void stop() {
    if (!running_.exchange(false)) return;
    char c = 1;
    (void)!::write(wake_[1], &c, 1);
    if (thread_.joinable()) thread_.join();
    for (auto& c2 : conns_) ::close(c2.fd);
    conns_.clear();
    if (listen_fd_ >= 0) ::close(listen_fd_);
    ::close(wake_[0]); ::close(wake_[1]);
  }
  int port() const { return port_; }

 private:
  void reply(int fd, int32_t st, const std::string& val, int64_t num) {
    std::vector<char> out(4 + 4 + val.size() + 8);
    uint32_t vl = (uint32_t)val.size();
    memcpy(out.data(), &st, 4);
    memcpy(out.data() + 4, &vl, 4);
    if (vl) memcpy(out.data() + 8, val.data(), vl);
    memcpy(out.data() + 8 + vl, &num, 8);
    send_all(fd, out.data(), out.size());
  }
  void handle(int fd, const Request& r) {
    switch (r.op) {
      case OP_SET:
        kv_[r.key] = r.val;
        reply(fd, ST_OK, "", 0);
        wake_waiters(r.key);
        break;
      case OP_GET: {
        auto it = kv_.find(r.key);
        if (it != kv_.end()) { reply(fd, ST_OK, it->second, 0); break; }
        Waiter w{fd, r.key, Clock::now() + std::chrono::milliseconds(r.arg < 0 ? 0 : r.arg), r.arg >= 0};
        if (r.arg == 0) reply(fd, ST_MISSING, "", 0); else waiters_.push_back(w);
        break;
      }
      case OP_ADD: {
        int64_t cur = 0;
        auto it = kv_.find(r.key);
        if (it != kv_.end() && it->second.size() == 8) memcpy(&cur, it->second.data(), 8);
        cur += r.arg;
        kv_[r.key] = std::string(reinterpret_cast<char*>(&cur), 8);
        reply(fd, ST_OK, "", cur);
        wake_waiters(r.key);
        break;
      }
      case OP_CHECK: reply(fd, kv_.count(r.key) ? ST_OK : ST_MISSING, "", 0); break;
      case OP_DELETE: reply(fd, kv_.erase(r.key) ? ST_OK : ST_MISSING, "", 0); break;
      case OP_NUMKEYS: reply(fd, ST_OK, "", (int64_t)kv_.size()); break;
      case OP_PING: reply(fd, ST_OK, "", r.arg); break;
      default: reply(fd, ST_BAD, "", 0);
    }
  }
  void wake_waiters(const std::string& key) {
    for (size_t i = 0; i < waiters_.size();) {
      if (waiters_[i].key == key) {
        reply(waiters_[i].fd, ST_OK, kv_[key], 0);
        waiters_.erase(waiters_.begin() + i);
      } else ++i;
    }
  }
  // Parse as many complete requests as the connection's buffer holds.
  void drain(Conn& c) {
    for (;;) {
      const size_t n = c.inbuf.size();
      if (n < 5) return;
      uint32_t kl; memcpy(&kl, c.inbuf.data() + 1, 4);
      if (n < 5 + (size_t)kl + 4) return;
      uint32_t vl; memcpy(&vl, c.inbuf.data() + 5 + kl, 4);
      const size_t total = 5 + (size_t)kl + 4 + vl + 8;
      if (n < total) return;
      Request r;
      r.op = (uint8_t)c.inbuf[0];
      r.key.assign(c.inbuf.data() + 5, kl);
      r.val.assign(c.inbuf.data() + 9 + kl, vl);
      memcpy(&r.arg, c.inbuf.data() + 9 + kl + vl, 8);
      c.inbuf.erase(c.inbuf.begin(), c.inbuf.begin() + total);
      handle(c.fd, r);
    }
  }
  void loop() {
    while (running_) {
      std::vector<pollfd> pfds;
      pfds.push_back({listen_fd_, POLLIN, 0});
      pfds.push_back({wake_[0], POLLIN, 0});
      for (auto& c : conns_) pfds.push_back({c.fd, POLLIN, 0});
      int timeout = 200;
      auto now = Clock::now();
      for (auto& w : waiters_)
        if (w.has_deadline) {
          auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(w.deadline - now).count();
          if (ms < 0) ms = 0;
          if (ms < timeout) timeout = (int)ms;
        }
      int rc = ::poll(pfds.data(), pfds.size(), timeout);
      if (!running_) break;
      now = Clock::now();
      for (size_t i = 0; i < waiters_.size();) {
        if (waiters_[i].has_deadline && waiters_[i].deadline <= now) {
          reply(waiters_[i].fd, ST_TIMEOUT, "", 0);
          waiters_.erase(waiters_.begin() + i);
        } else ++i;
      }
      if (rc <= 0) continue;
      if (pfds[0].revents & POLLIN) {
        int fd = ::accept(listen_fd_, nullptr, nullptr);
        if (fd >= 0) {
          int one = 1;
          setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
          conns_.push_back(Conn{fd, {}});
        }
      }
      const size_t nconn = pfds.size() - 2;      // conns present when pfds was built
      std::vector<int> dead;
      for (size_t i = 0; i < nconn; ++i) {
        if (!(pfds[i + 2].revents & (POLLIN | POLLHUP | POLLERR))) continue;
        Conn& c = conns_[i];
        char buf[65536];
        ssize_t k = ::recv(c.fd, buf, sizeof(buf), 0);
        if (k <= 0) { if (k < 0 && (errno == EINTR || errno == EAGAIN)) continue; dead.push_back(c.fd); continue; }
        c.inbuf.insert(c.inbuf.end(), buf, buf + k);
        drain(c);
      }
      for (int fd : dead) {
        for (size_t i = 0; i < waiters_.size();)
          if (waiters_[i].fd == fd) waiters_.erase(waiters_.begin() + i); else ++i;
        for (size_t i = 0; i < conns_.size(); ++i)
          if (conns_[i].fd == fd) { ::close(fd); conns_.erase(conns_.begin() + i); break; }
      }
    }
  }

  int listen_fd_ = -1, port_ = 0;
  int wake_[2] = {-1, -1};
  std::atomic<bool> running_{false};
  std::thread thread_;
  std::vector<Conn> conns_;
  std::vector<Waiter> waiters_;
  std::map<std::string, std::string> kv_;
};

struct Client {
  int fd = -1;
  std::mutex mu;
};

int request(Client* c, uint8_t op, const char* key, const void* val, uint32_t vlen, int64_t arg, int32_t* st,
            std::string* out, int64_t* num) {
  std::lock_guard<std::mutex> lk(c->mu);
  const uint32_t kl = (uint32_t)strlen(key);
  std::vector<char> buf(1 + 4 + kl + 4 + vlen + 8);
  buf[0] = (char)op;
  memcpy(buf.data() + 1, &kl, 4);
  memcpy(buf.data() + 5, key, kl);
  memcpy(buf.data() + 5 + kl, &vlen, 4);
  if (vlen) memcpy(buf.data() + 9 + kl, val, vlen);
  memcpy(buf.data() + 9 + kl + vlen, &arg, 8);
  if (!send_all(c->fd, buf.data(), buf.size())) return fail(-1, "store: connection to the master lost (send)");
  char hdr[8];
  if (!recv_all(c->fd, hdr, 8)) return fail(-1, "store: connection to the master lost (recv)");
  uint32_t vl;
  memcpy(st, hdr, 4);
  memcpy(&vl, hdr + 4, 4);
  std::string v(vl, '\0');
  if (vl && !recv_all(c->fd, &v[0], vl)) return fail(-1, "store: short read");
  int64_t n;
  if (!recv_all(c->fd, &n, 8)) return fail(-1, "store: short read");
  if (out) *out = std::move(v);
  if (num) *num = n;
  return 0;
}

}  // namespace

B200H_API const char* b200h_last_error() { return g_err; }

B200H_API void* b200h_store_server_start(const char* host, int port, int* actual_port) {
  Server* s = new Server();
  if (s->start(host, port) != 0) { delete s; return nullptr; }
  if (actual_port) *actual_port = s->port();
  return s;
}
B200H_API void b200h_store_server_stop(void* h) {
  if (!h) return;
  Server* s = static_cast<Server*>(h);
  s->stop();
  delete s;
}

// Retries until timeout_ms: the master may not be listening yet when a worker starts.
B200H_API void* b200h_store_connect(const char* host, int port, int timeout_ms) {
  const auto deadline = Clock::now() + std::chrono::milliseconds(timeout_ms);
  int last_errno = 0;
  for (;;) {
    int fd = ::socket(AF_INET, SOCK_STREAM, 0);
    if (fd < 0) { fail(-1, "store: socket() failed: %s", strerror(errno)); return nullptr; }
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    addr.sin_port = htons((uint16_t)port);
    if (inet_pton(AF_INET, host, &addr.sin_addr) != 1) {
      ::close(fd);
      fail(-1, "store: MASTER_ADDR '%s' is not an IPv4 address", host);
      return nullptr;
    }
    if (::connect(fd, (sockaddr*)&addr, sizeof(addr)) == 0) {
      int one = 1;
      setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
      Client* c = new Client();
      c->fd = fd;
      return c;
    }
    last_errno = errno;
    ::close(fd);
    if (Clock::now() >= deadline) break;
    std::this_thread::sleep_for(std::chrono::milliseconds(20));
  }
  fail(-1, "store: could not reach the master at %s:%d within %d ms: %s", host, port, timeout_ms,
       strerror(last_errno));
  return nullptr;
}
B200H_API void b200h_store_close(void* h) {
  if (!h) return;
  Client* c = static_cast<Client*>(h);
  ::close(c->fd);
  delete c;
}
B200H_API int b200h_store_set(void* h, const char* key, const void* val, int len) {
  int32_t st;
  int rc = request(static_cast<Client*>(h), OP_SET, key, val, (uint32_t)len, 0, &st, nullptr, nullptr);
  if (rc) return rc;
  return st == ST_OK ? 0 : fail(st, "store: set('%s') failed", key);
}
// timeout_ms < 0: wait forever; 0: do not wait.  Returns 0, or 1 on timeout / 2 if missing.
B200H_API int b200h_store_get(void* h, const char* key, void* buf, int cap, int* len_out, int timeout_ms) {
  int32_t st; std::string v;
  int rc = request(static_cast<Client*>(h), OP_GET, key, nullptr, 0, timeout_ms, &st, &v, nullptr);
  if (rc) return rc;
  if (st == ST_TIMEOUT) return fail(1, "store: timed out after %d ms waiting for key '%s'", timeout_ms, key);
  if (st != ST_OK) return fail(2, "store: key '%s' not found", key);
  if (len_out) *len_out = (int)v.size();
  if ((int)v.size() > cap) return fail(3, "store: value of '%s' (%zu B) exceeds buffer (%d B)", key, v.size(), cap);
  if (!v.empty()) memcpy(buf, v.data(), v.size());
  return 0;
}
B200H_API int b200h_store_add(void* h, const char* key, long long delta, long long* result) {
  int32_t st; int64_t n = 0;
  int rc = request(static_cast<Client*>(h), OP_ADD, key, nullptr, 0, delta, &st, nullptr, &n);
  if (rc) return rc;
  if (result) *result = n;
  return 0;
}
B200H_API int b200h_store_check(void* h, const char* key) {
  int32_t st;
  int rc = request(static_cast<Client*>(h), OP_CHECK, key, nullptr, 0, 0, &st, nullptr, nullptr);
  if (rc) return -1;
  return st == ST_OK ? 1 : 0;
}
B200H_API int b200h_store_delete(void* h, const char* key) {
  int32_t st;
  int rc = request(static_cast<Client*>(h), OP_DELETE, key, nullptr, 0, 0, &st, nullptr, nullptr);
  if (rc) return -1;
  return st == ST_OK ? 1 : 0;
}
B200H_API long long b200h_store_num_keys(void* h) {
  int32_t st; int64_t n = 0;
  int rc = request(static_cast<Client*>(h), OP_NUMKEYS, "", nullptr, 0, 0, &st, nullptr, &n);
  return rc ? -1 : n;
}
