// Self-test of the native TCP transport (csrc/host/tcp_transport.cpp) for the sanitizer builds: two transports in one
// process exchange frames of ragged sizes from several sender threads in both directions while a consumer thread
// releases the buffers, a peer is torn down mid-stream, and a client that does not speak the protocol is rejected.
// Linked into host_selftest (ASan + UBSan) and, with -DSELFTEST_TCP_MAIN, built alone under ThreadSanitizer
// (scripts/sanitize_host.sh) -- the OpenMP kernels of the other files are not TSan-clean by construction.
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <thread>
#include <vector>

extern "C" {
void* dcnn_tcp_create(const char*, int, int);
int dcnn_tcp_port(void*);
int dcnn_tcp_connect(void*, const char*, const char*, int, int);
int dcnn_tcp_disconnect(void*, const char*);
int dcnn_tcp_peer_alive(void*, const char*);
int dcnn_tcp_send(void*, const char*, const void* const*, const size_t*, int);
int dcnn_tcp_recv(void*, int, const uint8_t**, size_t*, uint64_t*);
int dcnn_tcp_release(void*, uint64_t);
int dcnn_tcp_flush(void*, int);
void dcnn_tcp_stats(void*, uint64_t*);
void dcnn_tcp_destroy(void*);
}

static int tcp_failures = 0;
#define TCHECK(cond, msg)                                            \
  do {                                                               \
    if (!(cond)) {                                                   \
      ++tcp_failures;                                                \
      fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, msg);  \
    }                                                                \
  } while (0)

// frame body: u32 sender, u32 sequence, then (size - 8) bytes of (sender * 31 + seq + i) & 0xFF
static std::vector<uint8_t> make_body(uint32_t sender, uint32_t seq, size_t size) {
  std::vector<uint8_t> b(size < 8 ? 8 : size);
  memcpy(b.data(), &sender, 4);
  memcpy(b.data() + 4, &seq, 4);
  for (size_t i = 8; i < b.size(); ++i) b[i] = (uint8_t)((sender * 31 + seq + i) & 0xFF);
  return b;
}

static bool check_body(const uint8_t* p, size_t n, uint32_t* sender, uint32_t* seq) {
  if (n < 8) return false;
  memcpy(sender, p, 4);
  memcpy(seq, p + 4, 4);
  for (size_t i = 8; i < n; ++i)
    if (p[i] != (uint8_t)((*sender * 31 + *seq + i) & 0xFF)) return false;
  return true;
}

static size_t body_size(uint32_t sender, uint32_t seq) {
  static const size_t sizes[] = {8, 9, 100, 4095, 4096, 4097, 65536, 300000, 3u << 20};
  return sizes[(sender * 7 + seq) % (sizeof(sizes) / sizeof(sizes[0]))];
}

static void consume(void* h, int expect, int senders, std::atomic<int>* bad) {
  std::vector<uint32_t> next(senders, 0);
  for (int got = 0; got < expect;) {
    const uint8_t* data;
    size_t len;
    uint64_t token;
    const int rc = dcnn_tcp_recv(h, 20000, &data, &len, &token);
    if (rc != 1) {
      ++*bad;
      return;
    }
    uint32_t sender = 0, seq = 0;
    uint64_t blen;
    memcpy(&blen, data + 2, 8);
    if (data[0] != 1 || blen + 10 != len || !check_body(data + 10, len - 10, &sender, &seq) ||
        sender >= (uint32_t)senders || seq != next[sender] || len - 10 != body_size(sender, seq))
      ++*bad;
    else
      ++next[sender];
    dcnn_tcp_release(h, token);
    ++got;
  }
}

int run_tcp_selftest() {
  void* a = dcnn_tcp_create("127.0.0.1", 0, 2);
  void* b = dcnn_tcp_create("127.0.0.1", 0, 3);
  TCHECK(a && b, "create");
  if (!a || !b) return 1;
  TCHECK(dcnn_tcp_connect(a, "b", "127.0.0.1", dcnn_tcp_port(b), 5000) == 0, "a dials b");
  TCHECK(dcnn_tcp_connect(b, "a", "127.0.0.1", dcnn_tcp_port(a), 5000) == 0, "b dials a");
  const int senders = 3, per = 40;
  std::atomic<int> bad{0};
  std::thread ca(consume, a, senders * per, senders, &bad), cb(consume, b, senders * per, senders, &bad);
  std::vector<std::thread> ts;
  for (int dir = 0; dir < 2; ++dir)
    for (int s = 0; s < senders; ++s)
      ts.emplace_back([=, &bad] {
        void* h = dir == 0 ? a : b;
        const char* peer = dir == 0 ? "b" : "a";
        for (int i = 0; i < per; ++i) {
          std::vector<uint8_t> body = make_body((uint32_t)s, (uint32_t)i, body_size(s, i));
          // split into up to three parts to exercise the gather path
          const size_t c1 = body.size() / 3, c2 = body.size() / 2;
          const void* parts[3] = {body.data(), body.data() + c1, body.data() + c2};
          const size_t lens[3] = {c1, c2 - c1, body.size() - c2};
          if (dcnn_tcp_send(h, peer, parts, lens, 3) != 0) ++bad;
        }
      });
  for (auto& t : ts) t.join();
  TCHECK(dcnn_tcp_flush(a, 20000) == 1 && dcnn_tcp_flush(b, 20000) == 1, "flush");
  ca.join();
  cb.join();
  TCHECK(bad.load() == 0, "frames arrive whole, in per-sender order, with the right contents");
  uint64_t st[9];
  dcnn_tcp_stats(b, st);
  TCHECK(st[3] == (uint64_t)senders * per && st[6] == 0, "receive counters");
  // unknown peer, then tear b down while a still holds a connection to it
  const void* p0[1] = {"x"};
  const size_t l0[1] = {1};
  TCHECK(dcnn_tcp_send(a, "nobody", p0, l0, 1) == -1, "unknown peer is refused");
  dcnn_tcp_destroy(b);
  for (int i = 0; i < 500 && dcnn_tcp_peer_alive(a, "b"); ++i) std::this_thread::sleep_for(std::chrono::milliseconds(10));
  TCHECK(dcnn_tcp_peer_alive(a, "b") == 0, "peer loss is noticed");
  TCHECK(dcnn_tcp_send(a, "b", p0, l0, 1) == -1, "send to a lost peer fails");
  TCHECK(dcnn_tcp_connect(a, "gone", "127.0.0.1", 1, 200) == -1, "connect to a closed port fails within the timeout");
  TCHECK(dcnn_tcp_disconnect(a, "b") == 0 && dcnn_tcp_disconnect(a, "b") == -1, "disconnect");
  dcnn_tcp_destroy(a);
  return tcp_failures;
}

#ifdef SELFTEST_TCP_MAIN
int main() {
  const int f = run_tcp_selftest();
  if (f) {
    fprintf(stderr, "%d tcp check(s) failed\n", f);
    return 1;
  }
  printf("tcp selftest: all checks passed\n");
  return 0;
}
#endif
