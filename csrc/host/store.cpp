// TCP key-value store used for rendezvous (rank 0 hosts, everybody connects).
//
// Role in the framework: replaces the library TCPStore that the reference reaches through
// `dist.init_process_group("gloo", rank=, world_size=)` with MASTER_ADDR/MASTER_PORT in the
// environment (ref:main.py:47-50).  It only carries bootstrap traffic: socket addresses for the
// CPU mesh, CUDA memory handles for the symmetric heap, and host-side barriers.
//
// Wire format (all little-endian):
//   request : u8 op | u32 klen | key bytes | u64 vlen | value bytes
//   response: i64 status_or_len | payload bytes (GET only)
// ops: 1 SET, 2 GET (blocks server-side until the key exists), 3 ADD (i64 delta, returns new
// value), 4 CHECK, 5 DELETE, 6 NUMKEYS.
#include <arpa/inet.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "host_util.h"

namespace dcp {

enum : uint8_t { OP_SET = 1, OP_GET = 2, OP_ADD = 3, OP_CHECK = 4, OP_DELETE = 5, OP_NUMKEYS = 6 };

struct Conn {
  int fd = -1;
  std::vector<uint8_t> in;   // bytes received, not yet parsed
  std::vector<uint8_t> out;  // bytes to send
  std::string waiting_key;   // non-empty: blocked GET
};

class StoreServer {
 public:
  StoreServer(const char* host, int port) {
    listen_fd_ = tcp_listen(host, port, &port_);
    if (listen_fd_ < 0) return;
    if (pipe(wake_) != 0) wake_[0] = wake_[1] = -1;
    running_ = true;
    thread_ = std::thread([this] { loop(); });
  }
  ~StoreServer() { stop(); }
  bool ok() const { return listen_fd_ >= 0; }
  int port() const { return port_; }
  void stop() {
    if (!running_.exchange(false)) return;
    if (wake_[1] >= 0) { char c = 1; (void)!write(wake_[1], &c, 1); }
    if (thread_.joinable()) thread_.join();
    for (auto& c : conns_) close(c.fd);
    conns_.clear();
    close(listen_fd_);
    if (wake_[0] >= 0) { close(wake_[0]); close(wake_[1]); }
  }

 private:
  void reply(Conn& c, int64_t status, const std::string* payload = nullptr) {
    const uint8_t* p = reinterpret_cast<const uint8_t*>(&status);
    c.out.insert(c.out.end(), p, p + 8);
    if (payload) c.out.insert(c.out.end(), payload->begin(), payload->end());
  }
  // Parse as many complete requests as are buffered.
  void parse(Conn& c) {
    for (;;) {
      if (!c.waiting_key.empty()) return;  // blocked on GET: keep later requests queued
      if (c.in.size() < 5) return;
      uint8_t op = c.in[0];
      uint32_t klen; memcpy(&klen, &c.in[1], 4);
      if (c.in.size() < 5 + (size_t)klen + 8) return;
      uint64_t vlen; memcpy(&vlen, &c.in[5 + klen], 8);
      size_t total = 5 + (size_t)klen + 8 + vlen;
      if (c.in.size() < total) return;
      std::string key(reinterpret_cast<char*>(&c.in[5]), klen);
      std::string val(reinterpret_cast<char*>(&c.in[5 + klen + 8]), vlen);
      c.in.erase(c.in.begin(), c.in.begin() + total);
      switch (op) {
        case OP_SET: kv_[key] = std::move(val); reply(c, 0); wake_waiters(key); break;
        case OP_GET: {
          auto it = kv_.find(key);
          if (it != kv_.end()) reply(c, (int64_t)it->second.size(), &it->second);
          else c.waiting_key = key;
          break;
        }
        case OP_ADD: {
          int64_t delta = 0; if (val.size() == 8) memcpy(&delta, val.data(), 8);
          int64_t cur = 0;
          auto it = kv_.find(key);
          if (it != kv_.end() && it->second.size() == 8) memcpy(&cur, it->second.data(), 8);
          cur += delta;
          kv_[key] = std::string(reinterpret_cast<char*>(&cur), 8);
          reply(c, cur); wake_waiters(key);
          break;
        }
        case OP_CHECK: reply(c, kv_.count(key) ? 1 : 0); break;
        case OP_DELETE: reply(c, (int64_t)kv_.erase(key)); break;
        case OP_NUMKEYS: reply(c, (int64_t)kv_.size()); break;
        default: reply(c, -2); break;
      }
    }
  }
  void wake_waiters(const std::string& key) {
    for (auto& c : conns_) {
      if (c.waiting_key == key) {
        c.waiting_key.clear();
        auto& v = kv_[key];
        reply(c, (int64_t)v.size(), &v);
        pending_parse_ = true;
      }
    }
  }
  void loop() {
    while (running_) {
      std::vector<pollfd> pfds;
      pfds.push_back({listen_fd_, POLLIN, 0});
      pfds.push_back({wake_[0], POLLIN, 0});
      for (auto& c : conns_) pfds.push_back({c.fd, (short)(POLLIN | (c.out.empty() ? 0 : POLLOUT)), 0});
      int r = poll(pfds.data(), pfds.size(), 200);
      if (r < 0 && errno != EINTR) break;
      if (!running_) break;
      if (pfds[0].revents & POLLIN) {
        int fd = accept(listen_fd_, nullptr, nullptr);
        if (fd >= 0) { set_nodelay(fd); set_nonblocking(fd); Conn c; c.fd = fd; conns_.push_back(std::move(c)); }
      }
      size_t n = pfds.size() - 2;
      std::vector<int> dead;
      for (size_t i = 0; i < n && i < conns_.size(); ++i) {
        Conn& c = conns_[i];
        short ev = pfds[i + 2].revents;
        if (ev & POLLIN) {
          uint8_t buf[65536];
          for (;;) {
            ssize_t k = recv(c.fd, buf, sizeof buf, 0);
            if (k > 0) { c.in.insert(c.in.end(), buf, buf + k); if (k < (ssize_t)sizeof buf) break; }
            else if (k == 0) { dead.push_back((int)i); break; }
            else { if (errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR) dead.push_back((int)i); break; }
          }
          parse(c);
        } else if (ev & (POLLHUP | POLLERR)) {
          dead.push_back((int)i);
        }
      }
      // a SET may have unblocked GETs whose connection has further buffered requests
      while (pending_parse_) { pending_parse_ = false; for (auto& c : conns_) parse(c); }
      for (auto& c : conns_) {
        while (!c.out.empty()) {
          ssize_t k = send(c.fd, c.out.data(), c.out.size(), MSG_NOSIGNAL);
          if (k > 0) c.out.erase(c.out.begin(), c.out.begin() + k);
          else break;
        }
      }
      for (auto it = dead.rbegin(); it != dead.rend(); ++it) {
        close(conns_[*it].fd);
        conns_.erase(conns_.begin() + *it);
      }
    }
  }

  int listen_fd_ = -1, port_ = 0;
  int wake_[2] = {-1, -1};
  std::atomic<bool> running_{false};
  bool pending_parse_ = false;
  std::thread thread_;
  std::vector<Conn> conns_;
  std::map<std::string, std::string> kv_;
};

class StoreClient {
 public:
  StoreClient(const char* host, int port, double timeout_s) { fd_ = tcp_connect_retry(host, port, timeout_s); }
  ~StoreClient() { if (fd_ >= 0) close(fd_); }
  bool ok() const { return fd_ >= 0; }

  // Returns status (>=0) or -1 on IO error / -3 on timeout.
  int64_t request(uint8_t op, const char* key, const void* val, uint64_t vlen, double timeout_s, bool want_payload) {
    std::lock_guard<std::mutex> g(mu_);
    uint32_t klen = (uint32_t)strlen(key);
    std::vector<uint8_t> msg(5 + klen + 8 + vlen);
    msg[0] = op; memcpy(&msg[1], &klen, 4); memcpy(&msg[5], key, klen);
    memcpy(&msg[5 + klen], &vlen, 8);
    if (vlen) memcpy(&msg[5 + klen + 8], val, vlen);
    if (!send_all(fd_, msg.data(), msg.size())) return -1;
    int64_t status;
    int rc = recv_all_timeout(fd_, &status, 8, timeout_s);
    if (rc != 0) { if (rc == 1) { poisoned_ = true; return -3; } return -1; }
    last_.clear();
    if (want_payload && status > 0) {
      last_.resize((size_t)status);
      if (recv_all_timeout(fd_, last_.data(), last_.size(), -1.0) != 0) return -1;
    }
    return status;
  }
  std::vector<uint8_t> last_;
  bool poisoned_ = false;  // a timed-out GET leaves a reply in flight; the connection is unusable after

 private:
  int fd_ = -1;
  std::mutex mu_;
};

}  // namespace dcp

using namespace dcp;

extern "C" {

void* dcp_store_server_start(const char* host, int port, int* out_port) {
  auto* s = new StoreServer(host, port);
  if (!s->ok()) { delete s; return nullptr; }
  if (out_port) *out_port = s->port();
  return s;
}
void dcp_store_server_stop(void* h) { delete static_cast<StoreServer*>(h); }

void* dcp_store_client_connect(const char* host, int port, double timeout_s) {
  auto* c = new StoreClient(host, port, timeout_s);
  if (!c->ok()) { delete c; return nullptr; }
  return c;
}
void dcp_store_client_close(void* h) { delete static_cast<StoreClient*>(h); }

int64_t dcp_store_set(void* h, const char* key, const void* val, int64_t len) {
  return static_cast<StoreClient*>(h)->request(OP_SET, key, val, (uint64_t)len, -1.0, false);
}
// Blocks until `key` exists or the timeout expires (-3).  The value stays inside the client until the
// next call; fetch it with dcp_store_copy_last.
int64_t dcp_store_get(void* h, const char* key, double timeout_s) {
  auto* c = static_cast<StoreClient*>(h);
  if (c->poisoned_) return -1;
  return c->request(OP_GET, key, nullptr, 0, timeout_s, true);
}
int64_t dcp_store_copy_last(void* h, void* buf, int64_t cap) {
  auto* c = static_cast<StoreClient*>(h);
  int64_t n = (int64_t)c->last_.size();
  if (n > cap) return -1;
  if (n) memcpy(buf, c->last_.data(), (size_t)n);
  return n;
}
int64_t dcp_store_add(void* h, const char* key, int64_t delta) {
  return static_cast<StoreClient*>(h)->request(OP_ADD, key, &delta, 8, -1.0, false);
}
int64_t dcp_store_check(void* h, const char* key) {
  return static_cast<StoreClient*>(h)->request(OP_CHECK, key, nullptr, 0, -1.0, false);
}
int64_t dcp_store_delete(void* h, const char* key) {
  return static_cast<StoreClient*>(h)->request(OP_DELETE, key, nullptr, 0, -1.0, false);
}
int64_t dcp_store_num_keys(void* h) {
  return static_cast<StoreClient*>(h)->request(OP_NUMKEYS, "", nullptr, 0, -1.0, false);
}

}  // extern "C"
