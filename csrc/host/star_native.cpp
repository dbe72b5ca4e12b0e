// Host-side transport of the TCP star (parallel/star.py): the reference's message path in native code.
//
// The reference moves every message through  pickle -> mgzip.compress(level 1, 12 threads, 1 MB blocks) ->
// struct.pack('>I', len) + payload -> blocking socket  (R:43-54, R:119-120) and reads it back with a
// `data += packet` loop (R:75-102).  Here, as plain C functions over file descriptors (no Python, no torch, no
// CUDA; built into lib/libb200star.so and called through ctypes with the GIL released):
//
//   * star_deflate / star_inflate: standard multi-member gzip, one member per block, compressed AND inflated
//     in parallel.  Every member carries its own size in an FEXTRA subfield ("BS", 4 bytes), so a reader can
//     find the member boundaries without inflating; readers that do not know the field skip it (RFC 1952),
//     so Python's gzip / zlib and mgzip read the stream, and streams from those are read here (sequentially
//     when the field is absent).
//   * star_send_frame / star_recv_frame: the '>I' length prefix + payload, written with one writev and read
//     straight into a buffer of the announced size, every wait bounded by poll() (the reference blocks forever).
//   * star_gather_frames: the root's gather, reading from ALL workers as their bytes arrive instead of
//     draining worker 0 first, then 1, ... (R:215-220).
//
// Return codes: 0 ok, STAR_TIMEOUT, STAR_CLOSED (orderly or reset), STAR_OSERR (errno is left set),
// STAR_TOOBIG (> 4 GiB - 1), STAR_BADDATA (not gzip / CRC or length mismatch), STAR_NOMEM.
#include <arpa/inet.h>
#include <errno.h>
#include <fcntl.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/uio.h>
#include <time.h>
#include <unistd.h>
#include <zlib.h>

#include <atomic>
#include <thread>
#include <vector>

extern "C" {

enum { STAR_OK = 0, STAR_TIMEOUT = -1, STAR_CLOSED = -2, STAR_OSERR = -3, STAR_TOOBIG = -4, STAR_BADDATA = -5, STAR_NOMEM = -6 };

void star_free(void* p) { free(p); }
int star_abi_version() { return 2; }

}  // extern "C"

namespace {

constexpr size_t kHeader = 10 + 2 + 8;          // fixed gzip header + XLEN + one 'B','S' subfield with 4 data bytes
constexpr size_t kTrailer = 8;                  // CRC32 + ISIZE

inline void put32(uint8_t* p, uint32_t v) { p[0] = v & 255; p[1] = (v >> 8) & 255; p[2] = (v >> 16) & 255; p[3] = v >> 24; }
inline uint32_t get32(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }

// One gzip member for in[0, n): header with the size field, raw deflate, trailer.  Returns the member size or 0.
size_t deflate_member(const uint8_t* in, size_t n, int level, int strategy, uint8_t* out, size_t cap) {
  z_stream zs;
  memset(&zs, 0, sizeof(zs));
  if (deflateInit2(&zs, level, Z_DEFLATED, -15, 8, strategy == 1 ? Z_RLE : Z_DEFAULT_STRATEGY) != Z_OK) return 0;
  zs.next_in = const_cast<Bytef*>(in);
  zs.avail_in = (uInt)n;
  zs.next_out = out + kHeader;
  zs.avail_out = (uInt)(cap - kHeader - kTrailer);
  const int rc = deflate(&zs, Z_FINISH);
  const size_t body = zs.total_out;
  deflateEnd(&zs);
  if (rc != Z_STREAM_END) return 0;
  const size_t total = kHeader + body + kTrailer;
  uint8_t* h = out;
  h[0] = 0x1f; h[1] = 0x8b; h[2] = 8; h[3] = 4;                 // FLG.FEXTRA
  put32(h + 4, 0);                                              // MTIME 0: identical input -> identical bytes
  h[8] = (level >= 9) ? 2 : (level <= 1 ? 4 : 0);
  h[9] = 255;                                                   // OS unknown
  h[10] = 8; h[11] = 0;                                         // XLEN
  h[12] = 'B'; h[13] = 'S'; h[14] = 4; h[15] = 0;
  put32(h + 16, (uint32_t)total);
  put32(out + kHeader + body, (uint32_t)crc32(crc32(0L, Z_NULL, 0), in, (uInt)n));
  put32(out + kHeader + body + 4, (uint32_t)n);
  return total;
}

struct Member { size_t off, size, body_off, out_off; uint32_t isize; };

// Walks the members using their size fields.  False if any member lacks the field (or the data is not gzip).
bool index_members(const uint8_t* in, size_t n, std::vector<Member>& ms) {
  size_t off = 0, out_off = 0;
  while (off < n) {
    if (n - off < 18 || in[off] != 0x1f || in[off + 1] != 0x8b || in[off + 2] != 8) return false;
    const uint8_t flg = in[off + 3];
    if (!(flg & 4) || (flg & ~4)) return false;                 // only FEXTRA, nothing else (name, comment, hcrc)
    const size_t xlen = in[off + 10] | ((size_t)in[off + 11] << 8);
    if (n - off < 12 + xlen + kTrailer) return false;
    size_t q = off + 12, end = off + 12 + xlen, size = 0;
    while (q + 4 <= end) {
      const size_t len = in[q + 2] | ((size_t)in[q + 3] << 8);
      if (q + 4 + len > end) return false;
      if (in[q] == 'B' && in[q + 1] == 'S' && len == 4) size = get32(in + q + 4);
      q += 4 + len;
    }
    if (size < 12 + xlen + kTrailer || size > n - off) return false;
    Member m;
    m.off = off; m.size = size; m.body_off = off + 12 + xlen; m.out_off = out_off;
    m.isize = get32(in + off + size - 4);
    ms.push_back(m);
    out_off += m.isize;
    off += size;
  }
  return true;
}

bool inflate_member(const uint8_t* in, const Member& m, uint8_t* out) {
  z_stream zs;
  memset(&zs, 0, sizeof(zs));
  if (inflateInit2(&zs, -15) != Z_OK) return false;
  zs.next_in = const_cast<Bytef*>(in + m.body_off);
  zs.avail_in = (uInt)(m.off + m.size - kTrailer - m.body_off);
  zs.next_out = out + m.out_off;
  zs.avail_out = m.isize;
  const int rc = inflate(&zs, Z_FINISH);
  const bool ok = (rc == Z_STREAM_END) && zs.total_out == m.isize;
  inflateEnd(&zs);
  if (!ok) return false;
  return (uint32_t)crc32(crc32(0L, Z_NULL, 0), out + m.out_off, m.isize) == get32(in + m.off + m.size - 8);
}

template <typename F>
void parallel_for(size_t count, int threads, F&& fn) {
  if (threads < 1) threads = 1;
  if ((size_t)threads > count) threads = (int)count;
  if (threads <= 1) {
    for (size_t i = 0; i < count; ++i) fn(i);
    return;
  }
  std::atomic<size_t> next{0};
  std::vector<std::thread> pool;
  pool.reserve(threads);
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&] {
      for (size_t i = next.fetch_add(1); i < count; i = next.fetch_add(1)) fn(i);
    });
  for (auto& th : pool) th.join();
}

// Any gzip stream (members without the size field, names, comments, ...): zlib's own gzip decoder, one member
// after the other, into a growing buffer.
int inflate_sequential(const uint8_t* in, size_t n, uint8_t** out, size_t* out_n) {
  if (n > 0xFFFFFFFFull) return STAR_TOOBIG;                      // a frame never is (the '>I' prefix)
  size_t cap = n * 4 + 65536, len = 0;
  uint8_t* buf = (uint8_t*)malloc(cap);
  if (!buf) return STAR_NOMEM;
  z_stream zs;
  memset(&zs, 0, sizeof(zs));
  if (inflateInit2(&zs, 31) != Z_OK) { free(buf); return STAR_NOMEM; }
  zs.next_in = const_cast<Bytef*>(in);
  zs.avail_in = (uInt)n;
  int result = STAR_OK;
  for (;;) {
    if (cap - len < 65536) {
      cap *= 2;
      uint8_t* nb = (uint8_t*)realloc(buf, cap);
      if (!nb) { result = STAR_NOMEM; break; }
      buf = nb;
    }
    const size_t room = (cap - len) > (1u << 30) ? (1u << 30) : (cap - len);
    zs.next_out = buf + len;
    zs.avail_out = (uInt)room;
    const int rc = inflate(&zs, Z_NO_FLUSH);
    len += room - zs.avail_out;
    if (rc == Z_STREAM_END) {                                   // end of one member: more may follow
      if (zs.avail_in == 0) break;
      if (inflateReset(&zs) != Z_OK) { result = STAR_BADDATA; break; }
    } else if (rc == Z_OK || (rc == Z_BUF_ERROR && zs.avail_out == 0)) {
      if (rc == Z_OK && zs.avail_in == 0 && zs.avail_out != 0) { result = STAR_BADDATA; break; }   // truncated
    } else {
      result = STAR_BADDATA;
      break;
    }
  }
  inflateEnd(&zs);
  if (result != STAR_OK) { free(buf); return result; }
  *out = buf;
  *out_n = len;
  return STAR_OK;
}

long long now_ms() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (long long)ts.tv_sec * 1000 + ts.tv_nsec / 1000000;
}

// poll one fd for `events` until the deadline (deadline < 0: forever).  0 ready, STAR_TIMEOUT, STAR_OSERR.
int wait_fd(int fd, short events, long long deadline) {
  for (;;) {
    int to = -1;
    if (deadline >= 0) {
      const long long left = deadline - now_ms();
      if (left <= 0) return STAR_TIMEOUT;
      to = left > 1000000 ? 1000000 : (int)left;
    }
    pollfd p{fd, events, 0};
    const int rc = poll(&p, 1, to);
    if (rc > 0) return STAR_OK;                                 // readable / writable / hung up: let the I/O call tell
    if (rc < 0 && errno != EINTR) return STAR_OSERR;
  }
}

inline bool closed_errno(int e) { return e == EPIPE || e == ECONNRESET || e == ENOTCONN || e == ECONNABORTED; }

int send_all(int fd, iovec* iov, int iovcnt, long long deadline) {
  while (iovcnt > 0) {
    msghdr mh;
    memset(&mh, 0, sizeof(mh));
    mh.msg_iov = iov;
    mh.msg_iovlen = iovcnt;
    const ssize_t k = sendmsg(fd, &mh, MSG_NOSIGNAL | MSG_DONTWAIT);
    if (k < 0) {
      if (errno == EINTR) continue;
      if (errno == EAGAIN || errno == EWOULDBLOCK) {
        const int w = wait_fd(fd, POLLOUT, deadline);
        if (w != STAR_OK) return w;
        continue;
      }
      return closed_errno(errno) ? STAR_CLOSED : STAR_OSERR;
    }
    size_t done = (size_t)k;
    while (iovcnt > 0 && done >= iov->iov_len) { done -= iov->iov_len; ++iov; --iovcnt; }
    if (iovcnt > 0) { iov->iov_base = (char*)iov->iov_base + done; iov->iov_len -= done; }
  }
  return STAR_OK;
}

// Reads up to n bytes without blocking.  >0 bytes read, 0 nothing available now, <0 error code.
ssize_t recv_some(int fd, uint8_t* p, size_t n) {
  for (;;) {
    const ssize_t k = recv(fd, p, n, MSG_DONTWAIT);
    if (k > 0) return k;
    if (k == 0) return STAR_CLOSED;
    if (errno == EINTR) continue;
    if (errno == EAGAIN || errno == EWOULDBLOCK) return 0;
    return closed_errno(errno) ? STAR_CLOSED : STAR_OSERR;
  }
}

int recv_exact(int fd, uint8_t* p, size_t n, long long deadline) {
  size_t got = 0;
  while (got < n) {
    const ssize_t k = recv_some(fd, p + got, n - got);
    if (k < 0) return (int)k;
    if (k == 0) {
      const int w = wait_fd(fd, POLLIN, deadline);
      if (w != STAR_OK) return w;
      continue;
    }
    got += (size_t)k;
  }
  return STAR_OK;
}

}  // namespace

extern "C" {

// ---- lossless stage (C2) ------------------------------------------------------------------------------------
// strategy: 0 zlib's default (what the reference's mgzip call does), 1 run-length matches only (Z_RLE): still plain
// deflate for any reader, and on a quantized gradient -- mostly runs of zeros -- both faster and smaller.
int star_deflate(const uint8_t* in, size_t n, int level, int strategy, int threads, size_t block, uint8_t** out,
                 size_t* out_n) {
  if (!out || !out_n || (n && !in)) return STAR_BADDATA;
  if (block < 1024) block = 1024;
  if (block > (1u << 30)) block = 1u << 30;
  if (level < 0 || level > 9) level = 1;
  const size_t count = n == 0 ? 1 : (n + block - 1) / block;
  std::vector<size_t> cap(count), off(count + 1, 0), size(count, 0);
  for (size_t i = 0; i < count; ++i) {
    const size_t len = (i + 1 == count) ? n - i * block : block;
    cap[i] = kHeader + kTrailer + compressBound((uLong)len) + 16;
    off[i + 1] = off[i] + cap[i];
  }
  uint8_t* scratch = (uint8_t*)malloc(off[count] ? off[count] : 1);
  if (!scratch) return STAR_NOMEM;
  std::atomic<bool> fail{false};
  parallel_for(count, threads, [&](size_t i) {
    const size_t len = (i + 1 == count) ? n - i * block : block;
    size[i] = deflate_member(in + i * block, len, level, strategy, scratch + off[i], cap[i]);
    if (size[i] == 0) fail = true;
  });
  if (fail) { free(scratch); return STAR_BADDATA; }
  size_t w = 0;                                                 // close the gaps between members, in place
  for (size_t i = 0; i < count; ++i) {
    if (w != off[i]) memmove(scratch + w, scratch + off[i], size[i]);
    w += size[i];
  }
  *out = scratch;
  *out_n = w;
  return STAR_OK;
}

int star_inflate(const uint8_t* in, size_t n, int threads, uint8_t** out, size_t* out_n) {
  if (!out || !out_n || (n && !in)) return STAR_BADDATA;
  if (n == 0) { *out = (uint8_t*)malloc(1); *out_n = 0; return *out ? STAR_OK : STAR_NOMEM; }
  std::vector<Member> ms;
  if (!index_members(in, n, ms)) return inflate_sequential(in, n, out, out_n);
  size_t total = 0;
  for (const Member& m : ms) total += m.isize;
  uint8_t* buf = (uint8_t*)malloc(total ? total : 1);
  if (!buf) return STAR_NOMEM;
  std::atomic<bool> fail{false};
  parallel_for(ms.size(), threads, [&](size_t i) {
    if (!inflate_member(in, ms[i], buf)) fail = true;
  });
  if (fail) { free(buf); return STAR_BADDATA; }
  *out = buf;
  *out_n = total;
  return STAR_OK;
}

// ---- framing over a connected socket (C3, C4) ------------------------------------------------------------
// The fd may be blocking or non-blocking (Python's settimeout() makes it non-blocking): every call is
// MSG_DONTWAIT + poll, bounded by timeout_ms (< 0: wait forever, as the reference does).
int star_send_frame(int fd, const uint8_t* buf, size_t n, int timeout_ms) {
  if (n > 0xFFFFFFFFull) return STAR_TOOBIG;
  uint8_t len[4] = {(uint8_t)(n >> 24), (uint8_t)(n >> 16), (uint8_t)(n >> 8), (uint8_t)n};        // '>I' (R:119)
  iovec iov[2] = {{len, 4}, {const_cast<uint8_t*>(buf), n}};
  return send_all(fd, iov, n ? 2 : 1, timeout_ms < 0 ? -1 : now_ms() + timeout_ms);
}

int star_recv_frame(int fd, uint8_t** out, size_t* out_n, int timeout_ms) {
  if (!out || !out_n) return STAR_BADDATA;
  const long long deadline = timeout_ms < 0 ? -1 : now_ms() + timeout_ms;
  uint8_t len[4];
  int rc = recv_exact(fd, len, 4, deadline);
  if (rc != STAR_OK) return rc;
  const size_t n = ((size_t)len[0] << 24) | ((size_t)len[1] << 16) | ((size_t)len[2] << 8) | len[3];
  uint8_t* buf = (uint8_t*)malloc(n ? n : 1);                   // sized once: no `data += packet` (R:93-101)
  if (!buf) return STAR_NOMEM;
  rc = recv_exact(fd, buf, n, deadline);
  if (rc != STAR_OK) { free(buf); return rc; }
  *out = buf;
  *out_n = n;
  return STAR_OK;
}

// The root's gather (C8): one frame from each of `count` sockets, read as the bytes arrive on ANY of them.
// outs[i] / lens[i] are filled for every socket that completed; on error the completed ones are freed.
int star_gather_frames(const int* fds, int count, uint8_t** outs, size_t* lens, int timeout_ms) {
  if (count < 0 || (count && (!fds || !outs || !lens))) return STAR_BADDATA;
  const long long deadline = timeout_ms < 0 ? -1 : now_ms() + timeout_ms;
  struct St { uint8_t hdr[4]; size_t got = 0, need = 0; uint8_t* buf = nullptr; bool have_len = false, done = false; };
  std::vector<St> st(count);
  std::vector<pollfd> pf(count);
  int remaining = count, err = STAR_OK;
  while (remaining > 0 && err == STAR_OK) {
    int to = -1;
    if (deadline >= 0) {
      const long long left = deadline - now_ms();
      if (left <= 0) { err = STAR_TIMEOUT; break; }
      to = left > 1000000 ? 1000000 : (int)left;
    }
    for (int i = 0; i < count; ++i) { pf[i].fd = st[i].done ? -1 : fds[i]; pf[i].events = POLLIN; pf[i].revents = 0; }
    const int rc = poll(pf.data(), (nfds_t)count, to);
    if (rc < 0) { if (errno == EINTR) continue; err = STAR_OSERR; break; }
    for (int i = 0; i < count && err == STAR_OK; ++i) {
      if (st[i].done || !pf[i].revents) continue;
      St& s = st[i];
      for (;;) {                                                // drain what is there
        uint8_t* dst = s.have_len ? s.buf + s.got : s.hdr + s.got;
        const size_t want = (s.have_len ? s.need : 4) - s.got;
        if (want == 0) {
          if (!s.have_len) {
            s.need = ((size_t)s.hdr[0] << 24) | ((size_t)s.hdr[1] << 16) | ((size_t)s.hdr[2] << 8) | s.hdr[3];
            s.buf = (uint8_t*)malloc(s.need ? s.need : 1);
            if (!s.buf) { err = STAR_NOMEM; break; }
            s.have_len = true;
            s.got = 0;
            continue;
          }
          s.done = true;
          --remaining;
          break;
        }
        const ssize_t k = recv_some(fds[i], dst, want);
        if (k < 0) { err = (int)k; break; }
        if (k == 0) break;
        s.got += (size_t)k;
      }
    }
  }
  for (int i = 0; i < count; ++i) {
    if (err == STAR_OK) { outs[i] = st[i].buf; lens[i] = st[i].need; }
    else { free(st[i].buf); outs[i] = nullptr; lens[i] = 0; }
  }
  return err;
}

// ---- rendezvous (C5, C6): plain TCP, TCP_NODELAY (R:175, R:200) -----------------------------------------
static int set_nodelay(int fd) {
  int one = 1;
  return setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
}

int star_listen(const char* addr, int port, int backlog, int* bound_port) {
  const int fd = socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
  if (fd < 0) return STAR_OSERR;
  int one = 1;
  setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  set_nodelay(fd);
  sockaddr_in sa;
  memset(&sa, 0, sizeof(sa));
  sa.sin_family = AF_INET;
  sa.sin_port = htons((uint16_t)port);
  if (!addr || !*addr || inet_pton(AF_INET, addr, &sa.sin_addr) != 1) sa.sin_addr.s_addr = htonl(INADDR_ANY);
  if (bind(fd, (sockaddr*)&sa, sizeof(sa)) < 0 || listen(fd, backlog > 0 ? backlog : 1) < 0) { close(fd); return STAR_OSERR; }
  if (bound_port) {
    socklen_t sl = sizeof(sa);
    if (getsockname(fd, (sockaddr*)&sa, &sl) == 0) *bound_port = ntohs(sa.sin_port);
  }
  return fd;
}

int star_accept(int lfd, int timeout_ms) {
  const int w = wait_fd(lfd, POLLIN, timeout_ms < 0 ? -1 : now_ms() + timeout_ms);
  if (w != STAR_OK) return w;
  const int fd = accept4(lfd, nullptr, nullptr, SOCK_CLOEXEC);
  if (fd < 0) return STAR_OSERR;
  set_nodelay(fd);
  return fd;
}

// Connects with retries (the reference connects once, R:201): the server may not be listening yet.
int star_connect(const char* addr, int port, int timeout_ms, int retries, int retry_delay_ms) {
  sockaddr_in sa;
  memset(&sa, 0, sizeof(sa));
  sa.sin_family = AF_INET;
  sa.sin_port = htons((uint16_t)port);
  if (!addr || inet_pton(AF_INET, addr, &sa.sin_addr) != 1) { errno = EINVAL; return STAR_OSERR; }
  for (int attempt = 0; attempt < (retries > 0 ? retries : 1); ++attempt) {
    const int fd = socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC | SOCK_NONBLOCK, 0);
    if (fd < 0) return STAR_OSERR;
    set_nodelay(fd);
    int rc = connect(fd, (sockaddr*)&sa, sizeof(sa));
    if (rc < 0 && errno == EINPROGRESS) {
      if (wait_fd(fd, POLLOUT, timeout_ms < 0 ? -1 : now_ms() + timeout_ms) == STAR_OK) {
        int soerr = 0;
        socklen_t sl = sizeof(soerr);
        getsockopt(fd, SOL_SOCKET, SO_ERROR, &soerr, &sl);
        rc = soerr == 0 ? 0 : -1;
        errno = soerr;
      }
    }
    if (rc == 0) {
      const int fl = fcntl(fd, F_GETFL, 0);
      fcntl(fd, F_SETFL, fl & ~O_NONBLOCK);
      return fd;
    }
    close(fd);
    timespec ts{retry_delay_ms / 1000, (long)(retry_delay_ms % 1000) * 1000000L};
    nanosleep(&ts, nullptr);
  }
  return STAR_TIMEOUT;
}

int star_close(int fd) { return close(fd) == 0 ? STAR_OK : STAR_OSERR; }

}  // extern "C"
