// Small blocking-socket helpers shared by the store, the CPU mesh and the FD-passing channel.
#pragma once
#include <arpa/inet.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <cerrno>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>

namespace dcp {

inline double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
inline void set_nodelay(int fd) { int one = 1; setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one); }
inline void set_nonblocking(int fd) { int fl = fcntl(fd, F_GETFL, 0); fcntl(fd, F_SETFL, fl | O_NONBLOCK); }

inline bool resolve(const char* host, int port, sockaddr_in* out) {
  memset(out, 0, sizeof *out);
  out->sin_family = AF_INET;
  out->sin_port = htons((uint16_t)port);
  if (!host || !*host || !strcmp(host, "localhost")) host = "127.0.0.1";
  if (inet_pton(AF_INET, host, &out->sin_addr) == 1) return true;
  addrinfo hints{}; hints.ai_family = AF_INET; hints.ai_socktype = SOCK_STREAM;
  addrinfo* res = nullptr;
  if (getaddrinfo(host, nullptr, &hints, &res) != 0 || !res) return false;
  out->sin_addr = reinterpret_cast<sockaddr_in*>(res->ai_addr)->sin_addr;
  freeaddrinfo(res);
  return true;
}

// Listen on host:port (port 0 = ephemeral).  Returns fd, or -1.
inline int tcp_listen(const char* host, int port, int* out_port) {
  int fd = socket(AF_INET, SOCK_STREAM, 0);
  if (fd < 0) return -1;
  int one = 1; setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
  sockaddr_in addr{};
  addr.sin_family = AF_INET; addr.sin_port = htons((uint16_t)port); addr.sin_addr.s_addr = htonl(INADDR_ANY);
  (void)host;  // bind on every interface; `host` only matters to the connecting side
  if (bind(fd, reinterpret_cast<sockaddr*>(&addr), sizeof addr) != 0 || listen(fd, 512) != 0) { close(fd); return -1; }
  socklen_t len = sizeof addr;
  getsockname(fd, reinterpret_cast<sockaddr*>(&addr), &len);
  if (out_port) *out_port = ntohs(addr.sin_port);
  return fd;
}

inline int tcp_connect_retry(const char* host, int port, double timeout_s) {
  sockaddr_in addr;
  if (!resolve(host, port, &addr)) return -1;
  double deadline = now_s() + (timeout_s > 0 ? timeout_s : 300.0);
  for (;;) {
    int fd = socket(AF_INET, SOCK_STREAM, 0);
    if (fd < 0) return -1;
    if (connect(fd, reinterpret_cast<sockaddr*>(&addr), sizeof addr) == 0) { set_nodelay(fd); return fd; }
    close(fd);
    if (now_s() > deadline) return -1;
    std::this_thread::sleep_for(std::chrono::milliseconds(20));
  }
}

inline bool send_all(int fd, const void* buf, size_t n) {
  const char* p = static_cast<const char*>(buf);
  while (n) {
    ssize_t k = send(fd, p, n, MSG_NOSIGNAL);
    if (k < 0) { if (errno == EINTR) continue; return false; }
    p += k; n -= (size_t)k;
  }
  return true;
}

// 0 ok, 1 timeout, -1 error/closed.  timeout_s < 0: wait forever.
inline int recv_all_timeout(int fd, void* buf, size_t n, double timeout_s) {
  char* p = static_cast<char*>(buf);
  double deadline = timeout_s >= 0 ? now_s() + timeout_s : 0;
  while (n) {
    if (timeout_s >= 0) {
      double left = deadline - now_s();
      if (left <= 0) return 1;
      pollfd pf{fd, POLLIN, 0};
      int r = poll(&pf, 1, (int)(left * 1000) + 1);
      if (r == 0) return 1;
      if (r < 0) { if (errno == EINTR) continue; return -1; }
    }
    ssize_t k = recv(fd, p, n, 0);
    if (k == 0) return -1;
    if (k < 0) { if (errno == EINTR) continue; return -1; }
    p += k; n -= (size_t)k;
  }
  return 0;
}

}  // namespace dcp
