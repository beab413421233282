// Full-mesh TCP transport between ranks: the CPU backend's wire (the role gloo's TCP pairs play for
// the reference, ref:main.py:50), plus Unix-socket file-descriptor passing that the symmetric-memory
// manager uses to hand cuMem POSIX handles to peers on the same host.
#include <sys/un.h>

#include <mutex>
#include <vector>

#include "host_util.h"

extern "C" {
int64_t dcp_store_set(void* h, const char* key, const void* val, int64_t len);
int64_t dcp_store_get(void* h, const char* key, double timeout_s);
int64_t dcp_store_copy_last(void* h, void* buf, int64_t cap);
}

namespace dcp {

struct MsgHeader { int64_t nbytes; int32_t tag; uint32_t magic; };
static constexpr uint32_t kMagic = 0xD15C0B20u;

struct Mesh {
  int rank = 0, world = 1;
  std::vector<int> fds;  // fds[peer], -1 for self
  double timeout_s = 1800.0;
  ~Mesh() { for (int fd : fds) if (fd >= 0) close(fd); }
};

static bool send_hdr(int fd, int64_t n, int tag) { MsgHeader h{n, tag, kMagic}; return send_all(fd, &h, sizeof h); }
static int recv_hdr(int fd, int64_t n, int tag, double timeout) {
  MsgHeader h;
  int rc = recv_all_timeout(fd, &h, sizeof h, timeout);
  if (rc != 0) return rc == 1 ? -3 : -1;
  if (h.magic != kMagic || h.nbytes != n || h.tag != tag) {
    fprintf(stderr, "[dcp mesh] message mismatch: got (bytes=%lld, tag=%d) expected (bytes=%lld, tag=%d)\n",
            (long long)h.nbytes, h.tag, (long long)n, tag);
    return -2;
  }
  return 0;
}

}  // namespace dcp
using namespace dcp;

extern "C" {

// Every rank listens on an ephemeral port and publishes "<prefix>/addr/<rank>" = "host:port".
// Rank i dials every j < i and accepts from every j > i; the dialler announces its rank.
void* dcp_mesh_create(void* store, int rank, int world, const char* my_host, const char* prefix, double timeout_s) {
  auto* m = new Mesh; m->rank = rank; m->world = world; m->fds.assign(world, -1);
  if (timeout_s > 0) m->timeout_s = timeout_s;
  int port = 0;
  int lfd = tcp_listen(my_host, 0, &port);
  if (lfd < 0) { delete m; return nullptr; }
  char key[256], val[256];
  snprintf(key, sizeof key, "%s/addr/%d", prefix, rank);
  snprintf(val, sizeof val, "%s:%d", (my_host && *my_host) ? my_host : "127.0.0.1", port);
  dcp_store_set(store, key, val, (int64_t)strlen(val));
  bool ok = true;
  for (int j = 0; j < rank && ok; ++j) {
    snprintf(key, sizeof key, "%s/addr/%d", prefix, j);
    int64_t n = dcp_store_get(store, key, m->timeout_s);
    if (n <= 0 || n >= (int64_t)sizeof val) { ok = false; break; }
    dcp_store_copy_last(store, val, sizeof val); val[n] = 0;
    char* colon = strrchr(val, ':');
    if (!colon) { ok = false; break; }
    *colon = 0;
    int fd = tcp_connect_retry(val, atoi(colon + 1), m->timeout_s);
    if (fd < 0) { ok = false; break; }
    int32_t me = rank;
    if (!send_all(fd, &me, 4)) { ok = false; close(fd); break; }
    m->fds[j] = fd;
  }
  for (int k = rank + 1; k < world && ok; ++k) {
    pollfd pf{lfd, POLLIN, 0};
    if (poll(&pf, 1, (int)(m->timeout_s * 1000)) <= 0) { ok = false; break; }
    int fd = accept(lfd, nullptr, nullptr);
    if (fd < 0) { ok = false; break; }
    set_nodelay(fd);
    int32_t who = -1;
    if (recv_all_timeout(fd, &who, 4, m->timeout_s) != 0 || who <= rank || who >= world || m->fds[who] >= 0) { ok = false; close(fd); break; }
    m->fds[who] = fd;
  }
  close(lfd);
  if (!ok) { delete m; return nullptr; }
  return m;
}

void dcp_mesh_destroy(void* h) { delete static_cast<Mesh*>(h); }

int dcp_mesh_send(void* h, int dst, const void* p, int64_t n, int tag) {
  auto* m = static_cast<Mesh*>(h);
  if (dst < 0 || dst >= m->world || dst == m->rank) return -4;
  int fd = m->fds[dst];
  if (!send_hdr(fd, n, tag)) return -1;
  return send_all(fd, p, (size_t)n) ? 0 : -1;
}

int dcp_mesh_recv(void* h, int src, void* p, int64_t n, int tag) {
  auto* m = static_cast<Mesh*>(h);
  if (src < 0 || src >= m->world || src == m->rank) return -4;
  int fd = m->fds[src];
  int rc = recv_hdr(fd, n, tag, m->timeout_s);
  if (rc) return rc;
  rc = recv_all_timeout(fd, p, (size_t)n, m->timeout_s);
  return rc == 0 ? 0 : (rc == 1 ? -3 : -1);
}

// Full-duplex exchange: progress the send to `dst` and the receive from `src` together so that a
// ring of ranks that all send first cannot deadlock on full socket buffers.
int dcp_mesh_sendrecv(void* h, int dst, const void* sp, int64_t sn, int src, void* rp, int64_t rn, int tag) {
  auto* m = static_cast<Mesh*>(h);
  if (dst < 0 || dst >= m->world || dst == m->rank || src < 0 || src >= m->world || src == m->rank) return -4;
  int sfd = m->fds[dst], rfd = m->fds[src];
  MsgHeader sh{sn, tag, kMagic}, rh{};
  // Linearised byte streams: header then payload, on both directions.
  size_t s_done = 0, r_done = 0;
  const size_t s_total = sizeof sh + (size_t)sn, r_total = sizeof rh + (size_t)rn;
  double deadline = now_s() + m->timeout_s;
  while (s_done < s_total || r_done < r_total) {
    pollfd pf[2]; int np = 0, si = -1, ri = -1;
    if (s_done < s_total) { pf[np] = {sfd, POLLOUT, 0}; si = np++; }
    if (r_done < r_total) { pf[np] = {rfd, POLLIN, 0}; ri = np++; }
    int r = poll(pf, np, 1000);
    if (r < 0 && errno != EINTR) return -1;
    if (now_s() > deadline) return -3;
    if (si >= 0 && (pf[si].revents & (POLLOUT | POLLERR | POLLHUP))) {
      const char* base; size_t off, len;
      if (s_done < sizeof sh) { base = reinterpret_cast<const char*>(&sh); off = s_done; len = sizeof sh - s_done; }
      else { base = static_cast<const char*>(sp); off = s_done - sizeof sh; len = s_total - s_done; }
      ssize_t k = send(sfd, base + off, len, MSG_NOSIGNAL | MSG_DONTWAIT);
      if (k > 0) s_done += (size_t)k;
      else if (k < 0 && errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR) return -1;
    }
    if (ri >= 0 && (pf[ri].revents & (POLLIN | POLLERR | POLLHUP))) {
      char* base; size_t off, len;
      if (r_done < sizeof rh) { base = reinterpret_cast<char*>(&rh); off = r_done; len = sizeof rh - r_done; }
      else { base = static_cast<char*>(rp); off = r_done - sizeof rh; len = r_total - r_done; }
      ssize_t k = recv(rfd, base + off, len, MSG_DONTWAIT);
      if (k > 0) {
        r_done += (size_t)k;
        if (r_done == sizeof rh && (rh.magic != kMagic || rh.nbytes != rn || rh.tag != tag)) {
          fprintf(stderr, "[dcp mesh] sendrecv mismatch: got (bytes=%lld, tag=%d) expected (bytes=%lld, tag=%d)\n",
                  (long long)rh.nbytes, rh.tag, (long long)rn, tag);
          return -2;
        }
      } else if (k == 0) return -1;
      else if (errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR) return -1;
    }
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------
// File-descriptor passing over abstract Unix sockets (SCM_RIGHTS).  Used to give peers on the same
// host a handle to cuMem allocations (POSIX FD shareable handles) without NCCL or NVSHMEM.
// ---------------------------------------------------------------------------------------------
static void make_abstract(sockaddr_un* a, socklen_t* len, const char* name) {
  memset(a, 0, sizeof *a);
  a->sun_family = AF_UNIX;
  size_t n = strlen(name);
  if (n > sizeof a->sun_path - 2) n = sizeof a->sun_path - 2;
  memcpy(a->sun_path + 1, name, n);  // leading NUL = abstract namespace, vanishes with the process
  *len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
}

int dcp_fd_listen(const char* name) {
  int fd = socket(AF_UNIX, SOCK_STREAM, 0);
  if (fd < 0) return -1;
  sockaddr_un a; socklen_t len; make_abstract(&a, &len, name);
  if (bind(fd, reinterpret_cast<sockaddr*>(&a), len) != 0 || listen(fd, 128) != 0) { close(fd); return -1; }
  return fd;
}

// Connect to `name` and send `nfds` descriptors tagged with the sender's rank.
int dcp_fd_send(const char* name, int my_rank, const int* fds, int nfds, double timeout_s) {
  double deadline = now_s() + (timeout_s > 0 ? timeout_s : 300.0);
  int s = -1;
  for (;;) {
    s = socket(AF_UNIX, SOCK_STREAM, 0);
    if (s < 0) return -1;
    sockaddr_un a; socklen_t len; make_abstract(&a, &len, name);
    if (connect(s, reinterpret_cast<sockaddr*>(&a), len) == 0) break;
    close(s);
    if (now_s() > deadline) return -3;
    std::this_thread::sleep_for(std::chrono::milliseconds(10));
  }
  int32_t hdr[2] = {my_rank, nfds};
  iovec iov{hdr, sizeof hdr};
  char cbuf[CMSG_SPACE(sizeof(int) * 16)];
  if (nfds > 16) { close(s); return -4; }
  memset(cbuf, 0, sizeof cbuf);
  msghdr msg{};
  msg.msg_iov = &iov; msg.msg_iovlen = 1;
  msg.msg_control = cbuf; msg.msg_controllen = CMSG_SPACE(sizeof(int) * nfds);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET; c->cmsg_type = SCM_RIGHTS; c->cmsg_len = CMSG_LEN(sizeof(int) * nfds);
  memcpy(CMSG_DATA(c), fds, sizeof(int) * nfds);
  int rc = sendmsg(s, &msg, MSG_NOSIGNAL) < 0 ? -1 : 0;
  // wait for the receiver's ack so the descriptors are installed before we close ours
  char ack = 0;
  if (rc == 0 && recv_all_timeout(s, &ack, 1, timeout_s > 0 ? timeout_s : 300.0) != 0) rc = -1;
  close(s);
  return rc;
}

// Accept one connection on `lfd`, receive its descriptors.  Returns the sender's rank (>=0) or <0.
int dcp_fd_recv(int lfd, int* out_fds, int max_fds, int* out_nfds, double timeout_s) {
  pollfd pf{lfd, POLLIN, 0};
  if (poll(&pf, 1, (int)((timeout_s > 0 ? timeout_s : 300.0) * 1000)) <= 0) return -3;
  int s = accept(lfd, nullptr, nullptr);
  if (s < 0) return -1;
  int32_t hdr[2] = {-1, 0};
  iovec iov{hdr, sizeof hdr};
  char cbuf[CMSG_SPACE(sizeof(int) * 16)];
  msghdr msg{};
  msg.msg_iov = &iov; msg.msg_iovlen = 1; msg.msg_control = cbuf; msg.msg_controllen = sizeof cbuf;
  ssize_t k = recvmsg(s, &msg, 0);
  int rc = -1;
  if (k == (ssize_t)sizeof hdr) {
    cmsghdr* c = CMSG_FIRSTHDR(&msg);
    if (c && c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) {
      int n = (int)((c->cmsg_len - CMSG_LEN(0)) / sizeof(int));
      if (n <= max_fds && n == hdr[1]) { memcpy(out_fds, CMSG_DATA(c), sizeof(int) * n); *out_nfds = n; rc = hdr[0]; }
    }
  }
  char ack = 1; send_all(s, &ack, 1);
  close(s);
  return rc;
}

void dcp_fd_close(int fd) { if (fd >= 0) close(fd); }

}  // extern "C"
