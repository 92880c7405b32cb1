// Sanitizer self-test of the native host runtime (libdcnn_host sources compiled straight into this executable with
// -fsanitize=address,undefined): every C-ABI entry point is driven over ragged / edge-case sizes and checked against
// naive loops, so out-of-bounds accesses, misaligned or overflowing arithmetic and leaks in the native code surface
// here rather than in a training run.  The reference advertises "debug builds with sanitizers" but never sets
// -fsanitize.  Built and run by tests/test_host_sanitizers.py and scripts/sanitize_host.sh.
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

extern "C" {
void dcnn_cpu_sgemm(int, int, int, int, int, float, const float*, int, const float*, int, float, float*, int);
void dcnn_cpu_dgemm(int, int, int, int, int, double, const double*, int, const double*, int, double, double*, int);
float dcnn_cpu_sdot(const float*, const float*, long long);
double dcnn_cpu_ssum(const float*, long long);
void dcnn_cpu_sbinary(int, const float*, const float*, float*, long long);
void dcnn_cpu_sunary(int, const float*, float*, long long);
void dcnn_cpu_stranspose(const float*, float*, int, int);
void dcnn_cpu_im2col(const float*, float*, int, int, int, int, int, int, int, int, int, int, int, int);
void dcnn_cpu_col2im(const float*, float*, int, int, int, int, int, int, int, int, int, int, int, int);
void dcnn_aug_flip(float*, int, int, int, int, const uint8_t*, int);
void dcnn_aug_affine_clamp(float*, int, size_t, const uint8_t*, const float*, const float*);
void dcnn_aug_random_crop(float*, int, int, int, int, int, const uint8_t*, const int*, const int*);
void dcnn_aug_cutout(float*, int, int, int, int, int, const uint8_t*, const int*, const int*);
void dcnn_aug_normalize(float*, int, int, int, int, const float*, const float*);
long long dcnn_encode_message(uint8_t*, size_t, const char*, const char*, uint16_t, uint64_t, uint64_t, const uint64_t*,
                              uint64_t, const float*, const uint8_t*, uint64_t, uint8_t);
int dcnn_decode_message(const uint8_t*, size_t, uint64_t*, uint64_t*, uint64_t*, uint64_t*, uint64_t*, uint16_t*,
                        uint64_t*, uint64_t*, uint64_t*, uint64_t*, uint64_t*, uint64_t*, uint8_t*);
long long dcnn_encode_tensor_record(uint8_t*, size_t, const uint64_t*, const float*);
long long dcnn_load_mnist_csv(const char*, float*, int*, long long, int);
void dcnn_cpu_bn_stats(const float*, int, int, int, float*, float*);
void dcnn_cpu_bn_apply(const float*, const float*, const float*, const float*, float*, int, int, int, int);
void dcnn_cpu_maxpool_fwd(const float*, float*, long long*, int, int, int, int, int, int, int, int, int, int, int, int);
void dcnn_cpu_maxpool_bwd(const float*, const long long*, float*, int, int, int, int);
void dcnn_cpu_avgpool_fwd(const float*, float*, int, int, int, int, int, int, int, int, int, int, int, int);
void dcnn_cpu_avgpool_bwd(const float*, float*, int, int, int, int, int, int, int, int, int, int, int, int);
void dcnn_cpu_conv2d_fwd(const float*, const float*, const float*, float*, int, int, int, int, int, int, int, int, int, int,
                         int, int, int);
void dcnn_cpu_conv2d_bwd(const float*, const float*, const float*, float*, float*, int, int, int, int, int, int, int, int,
                         int, int, int, int, int);
}

int run_tcp_selftest();   // selftest_tcp.cpp

static int failures = 0;
#define CHECK(cond, ...)                                  \
  do {                                                    \
    if (!(cond)) {                                        \
      ++failures;                                         \
      fprintf(stderr, "FAIL %s:%d: ", __FILE__, __LINE__); \
      fprintf(stderr, __VA_ARGS__);                       \
      fprintf(stderr, "\n");                              \
    }                                                     \
  } while (0)

static uint32_t rng_state = 12345u;
static float frand() {
  rng_state = rng_state * 1664525u + 1013904223u;
  return ((rng_state >> 8) & 0xFFFF) / 65536.0f - 0.5f;
}
static std::vector<float> randv(size_t n) {
  std::vector<float> v(n);
  for (auto& x : v) x = frand();
  return v;
}

static void test_gemm() {
  const int sizes[][3] = {{1, 1, 1}, {5, 7, 3}, {6, 16, 8}, {7, 17, 9}, {97, 33, 129}, {13, 100, 300}, {130, 5, 257}, {100, 130, 70}, {61, 97, 300}};
  for (auto& s : sizes) {
    const int M = s[0], N = s[1], K = s[2];
    for (int ta = 0; ta < 2; ++ta)
      for (int tb = 0; tb < 2; ++tb) {
        // exact-size allocations so that any overrun is caught by ASan's red zones
        std::vector<float> A = randv((size_t)M * K), B = randv((size_t)K * N), C = randv((size_t)M * N), R = C;
        const int lda = ta ? M : K, ldb = tb ? K : N;
        for (int i = 0; i < M; ++i)
          for (int j = 0; j < N; ++j) {
            double acc = 0;
            for (int k = 0; k < K; ++k)
              acc += (double)(ta ? A[(size_t)k * lda + i] : A[(size_t)i * lda + k]) *
                     (double)(tb ? B[(size_t)j * ldb + k] : B[(size_t)k * ldb + j]);
            R[(size_t)i * N + j] = (float)(0.75 * acc + 0.5 * R[(size_t)i * N + j]);
          }
        dcnn_cpu_sgemm(ta, tb, M, N, K, 0.75f, A.data(), lda, B.data(), ldb, 0.5f, C.data(), N);
        float worst = 0;
        for (size_t i = 0; i < C.size(); ++i) worst = fmaxf(worst, fabsf(C[i] - R[i]));
        CHECK(worst < 1e-3f, "sgemm %dx%dx%d ta=%d tb=%d: max abs err %g", M, N, K, ta, tb, worst);
      }
  }
  std::vector<double> A(35 * 20, 0.5), B(20 * 9, 2.0), C(35 * 9, 1e300);
  dcnn_cpu_dgemm(0, 0, 35, 9, 20, 1.0, A.data(), 20, B.data(), 9, 0.0, C.data(), 9);
  for (double v : C) CHECK(fabs(v - 20.0) < 1e-12, "dgemm value %g", v);
}

static void test_elementwise() {
  for (long long n : {0LL, 1LL, 7LL, 8LL, 33LL, 100003LL}) {
    std::vector<float> a = randv((size_t)n), b = randv((size_t)n), y((size_t)n);
    dcnn_cpu_sbinary(0, a.data(), b.data(), y.data(), n);
    for (long long i = 0; i < n; ++i) CHECK(y[i] == a[i] + b[i], "add mismatch at %lld", i);
    dcnn_cpu_sunary(8, a.data(), y.data(), n);
    for (long long i = 0; i < n; ++i) CHECK(y[i] == (a[i] > 0 ? a[i] : 0.f), "relu mismatch at %lld", i);
    double want = 0, s = 0;
    for (long long i = 0; i < n; ++i) { want += (double)a[i] * b[i]; s += a[i]; }
    CHECK(fabs(dcnn_cpu_sdot(a.data(), b.data(), n) - want) < 1e-2, "dot n=%lld", n);
    CHECK(fabs(dcnn_cpu_ssum(a.data(), n) - s) < 1e-3, "sum n=%lld", n);
  }
  std::vector<float> t = randv(37 * 41), o(37 * 41);
  dcnn_cpu_stranspose(t.data(), o.data(), 37, 41);
  for (int r = 0; r < 37; ++r)
    for (int c = 0; c < 41; ++c) CHECK(o[(size_t)c * 37 + r] == t[(size_t)r * 41 + c], "transpose");
}

static void test_im2col() {
  const int N = 2, C = 3, H = 7, W = 5, k = 3, s = 2, p = 1, OH = (H + 2 * p - k) / s + 1, OW = (W + 2 * p - k) / s + 1;
  std::vector<float> x = randv((size_t)N * C * H * W), col((size_t)C * k * k * N * OH * OW), back(x.size());
  dcnn_cpu_im2col(x.data(), col.data(), N, C, H, W, k, k, s, s, p, p, OH, OW);
  const size_t L = (size_t)N * OH * OW;
  std::vector<float> acc(x.size(), 0.f);
  for (int c = 0; c < C; ++c)
    for (int r = 0; r < k; ++r)
      for (int q = 0; q < k; ++q)
        for (int n = 0; n < N; ++n)
          for (int oh = 0; oh < OH; ++oh)
            for (int ow = 0; ow < OW; ++ow) {
              const int ih = oh * s - p + r, iw = ow * s - p + q;
              const float v = col[((size_t)(c * k + r) * k + q) * L + ((size_t)n * OH + oh) * OW + ow];
              const bool in = ih >= 0 && ih < H && iw >= 0 && iw < W;
              CHECK(v == (in ? x[(((size_t)n * C + c) * H + ih) * W + iw] : 0.f), "im2col value");
              if (in) acc[(((size_t)n * C + c) * H + ih) * W + iw] += v;
            }
  dcnn_cpu_col2im(col.data(), back.data(), N, C, H, W, k, k, s, s, p, p, OH, OW);
  for (size_t i = 0; i < back.size(); ++i) CHECK(fabsf(back[i] - acc[i]) < 1e-5f, "col2im at %zu", i);
}

static void test_augment() {
  const int n = 3, c = 2, h = 5, w = 4;
  std::vector<float> x = randv((size_t)n * c * h * w);
  for (auto& v : x) v += 0.5f;
  std::vector<float> y = x;
  const uint8_t mask[3] = {1, 0, 1};
  dcnn_aug_flip(y.data(), n, c, h, w, mask, 0);
  dcnn_aug_flip(y.data(), n, c, h, w, mask, 0);
  CHECK(y == x, "horizontal flip twice is not the identity");
  dcnn_aug_flip(y.data(), n, c, h, w, mask, 1);
  dcnn_aug_flip(y.data(), n, c, h, w, mask, 1);
  CHECK(y == x, "vertical flip twice is not the identity");
  const float mul[3] = {1.f, 1.f, 1.f}, add[3] = {10.f, 10.f, -10.f};
  dcnn_aug_affine_clamp(y.data(), n, (size_t)c * h * w, mask, mul, add);
  CHECK(y[0] == 1.f && y[(size_t)c * h * w] == x[(size_t)c * h * w] && y[(size_t)2 * c * h * w] == 0.f, "affine clamp");
  y = x;
  const int sx[3] = {0, 3, 8}, sy[3] = {8, 0, 4};  // pad 4: offsets in [0, 8]; (4,4) is the identity
  dcnn_aug_random_crop(y.data(), n, c, h, w, 4, mask, sx, sy);
  y = x;
  const int id[3] = {4, 4, 4};
  const uint8_t all[3] = {1, 1, 1};
  dcnn_aug_random_crop(y.data(), n, c, h, w, 4, all, id, id);
  CHECK(y == x, "crop at the padding offset is not the identity");
  const int cx[3] = {0, 3, 1}, cy[3] = {0, 4, 2};
  dcnn_aug_cutout(y.data(), n, c, h, w, 3, all, cx, cy);
  const float mean[2] = {0.5f, 0.25f}, sd[2] = {0.5f, 2.f};
  dcnn_aug_normalize(y.data(), n, c, h, w, mean, sd);
}

static void test_codec() {
  const uint64_t dims[4] = {2, 3, 4, 5};
  std::vector<float> t = randv(120);
  const long long need = dcnn_encode_message(nullptr, 0, "stage1", "coordinator", 7, 1, 42, dims, 4, t.data(), nullptr, 0, 0);
  CHECK(need > 0, "size query");
  std::vector<uint8_t> buf((size_t)need);
  CHECK(dcnn_encode_message(buf.data(), buf.size(), "stage1", "coordinator", 7, 1, 42, dims, 4, t.data(), nullptr, 0, 0) ==
            need, "encode");
  CHECK(dcnn_encode_message(buf.data(), buf.size() - 1, "stage1", "coordinator", 7, 1, 42, dims, 4, t.data(), nullptr, 0,
                            0) == -1, "short buffer must be refused");
  uint64_t bl, ro, rl, so, sl, pt, mb, d[8], nd, doff, dlen;
  uint16_t cmd;
  uint8_t flag;
  CHECK(dcnn_decode_message(buf.data(), buf.size(), &bl, &ro, &rl, &so, &sl, &cmd, &pt, &mb, d, &nd, &doff, &dlen, &flag) == 0,
        "decode");
  CHECK(cmd == 7 && pt == 1 && mb == 42 && nd == 4 && dlen == 480 && rl == 6 && sl == 11, "decoded header fields");
  CHECK(memcmp(buf.data() + doff, t.data(), 480) == 0, "payload bytes");
  // every truncation of a valid message must be rejected without reading past the end
  for (size_t cut = 0; cut < buf.size(); cut += 7) {
    std::vector<uint8_t> part(buf.begin(), buf.begin() + cut);
    CHECK(dcnn_decode_message(part.data(), part.size(), &bl, &ro, &rl, &so, &sl, &cmd, &pt, &mb, d, &nd, &doff, &dlen,
                              &flag) == -1, "truncated message (%zu bytes) accepted", cut);
  }
  const uint64_t shape[4] = {1, 2, 3, 4};
  std::vector<uint8_t> rec(32 + 96);
  CHECK(dcnn_encode_tensor_record(rec.data(), rec.size(), shape, t.data()) == 128, "tensor record");
  CHECK(dcnn_encode_tensor_record(rec.data(), rec.size() - 1, shape, t.data()) == -1, "short record buffer");
}

static void test_loader(const char* tmpdir) {
  const std::string path = std::string(tmpdir) + "/mini_mnist.csv";
  FILE* f = fopen(path.c_str(), "w");
  if (!f) { CHECK(false, "cannot write %s", path.c_str()); return; }
  fprintf(f, "label");
  for (int i = 0; i < 784; ++i) fprintf(f, ",p%d", i);
  fprintf(f, "\n");
  for (int r = 0; r < 3; ++r) {
    fprintf(f, "%d", r + 1);
    for (int i = 0; i < 784; ++i) fprintf(f, ",%d", (i + r) % 256);
    fprintf(f, "\n");
  }
  fclose(f);
  std::vector<float> img(3 * 784);
  std::vector<int> lab(3);
  const long long rows = dcnn_load_mnist_csv(path.c_str(), img.data(), lab.data(), 3, 784);
  CHECK(rows == 3 && lab[2] == 3, "mnist rows %lld", rows);
  CHECK(fabsf(img[784 + 5] - 6.f / 255.f) < 1e-6f, "pixel scaling");
  remove(path.c_str());
}

static void test_layers() {
  const int N = 2, C = 3, H = 7, W = 5, HW = H * W;
  std::vector<float> x = randv((size_t)N * C * HW), mean(C), var(C), y(x.size()), sc(C, 2.f), sh(C, -1.f);
  dcnn_cpu_bn_stats(x.data(), N, C, HW, mean.data(), var.data());
  for (int c = 0; c < C; ++c) {
    double s = 0;
    for (int n = 0; n < N; ++n)
      for (int i = 0; i < HW; ++i) s += x[((size_t)n * C + c) * HW + i];
    CHECK(fabs(mean[c] - s / (N * HW)) < 1e-5, "bn mean");
    CHECK(var[c] >= 0.f, "bn var");
  }
  dcnn_cpu_bn_apply(x.data(), sc.data(), sh.data(), nullptr, y.data(), N, C, HW, 1);
  for (size_t i = 0; i < y.size(); ++i) CHECK(y[i] == fmaxf(x[i] * 2.f - 1.f, 0.f), "bn apply");
  const int k = 3, st = 2, p = 1, OH = (H + 2 * p - k) / st + 1, OW = (W + 2 * p - k) / st + 1;
  std::vector<float> py((size_t)N * C * OH * OW), dx(x.size());
  std::vector<long long> idx(py.size());
  dcnn_cpu_maxpool_fwd(x.data(), py.data(), idx.data(), N, C, H, W, OH, OW, k, k, st, st, p, p);
  for (size_t i = 0; i < idx.size(); ++i) CHECK(idx[i] >= 0 && idx[i] < HW, "maxpool index in range");
  dcnn_cpu_maxpool_bwd(py.data(), idx.data(), dx.data(), N, C, HW, OH * OW);
  dcnn_cpu_avgpool_fwd(x.data(), py.data(), N, C, H, W, OH, OW, k, k, st, st, p, p);
  dcnn_cpu_avgpool_bwd(py.data(), dx.data(), N, C, H, W, OH, OW, k, k, st, st, p, p);
}

// blocked convolution against a naive loop nest, over shapes whose taps fall on padding, strides that skip pixels,
// kernels as large as the image and ragged last blocks
static void test_conv() {
  const int cases[][8] = {{5, 3, 9, 11, 7, 3, 2, 1}, {3, 4, 6, 6, 5, 3, 1, 2}, {2, 2, 5, 5, 3, 5, 1, 2}, {9, 3, 7, 7, 4, 3, 3, 0},
                          {2, 1, 1, 1, 2, 5, 2, 2}, {70, 2, 40, 40, 4, 3, 1, 1}, {33, 1, 28, 28, 8, 5, 1, 0}};
  for (const auto& cs : cases) {
    const int N = cs[0], C = cs[1], H = cs[2], W = cs[3], OC = cs[4], k = cs[5], st = cs[6], pd = cs[7];
    const int OH = (H + 2 * pd - k) / st + 1, OW = (W + 2 * pd - k) / st + 1;
    auto x = randv((size_t)N * C * H * W), w = randv((size_t)OC * C * k * k), b = randv(OC);
    auto dy = randv((size_t)N * OC * OH * OW);
    std::vector<float> y((size_t)N * OC * OH * OW), dx((size_t)N * C * H * W, 7.f), dw((size_t)OC * C * k * k, 0.25f);
    std::vector<double> yr(y.size()), dxr(dx.size(), 0.0), dwr(dw.size(), 0.25);
    for (int n = 0; n < N; ++n)
      for (int oc = 0; oc < OC; ++oc)
        for (int oh = 0; oh < OH; ++oh)
          for (int ow = 0; ow < OW; ++ow) {
            const size_t yo = (((size_t)n * OC + oc) * OH + oh) * OW + ow;
            double acc = b[oc];
            for (int c = 0; c < C; ++c)
              for (int r = 0; r < k; ++r)
                for (int s = 0; s < k; ++s) {
                  const int ih = oh * st - pd + r, iw = ow * st - pd + s;
                  if (ih < 0 || ih >= H || iw < 0 || iw >= W) continue;
                  const size_t xo = (((size_t)n * C + c) * H + ih) * W + iw, wo = (((size_t)oc * C + c) * k + r) * k + s;
                  acc += (double)x[xo] * w[wo];
                  dxr[xo] += (double)dy[yo] * w[wo];
                  dwr[wo] += (double)dy[yo] * x[xo];
                }
            yr[yo] = acc;
          }
    dcnn_cpu_conv2d_fwd(x.data(), w.data(), b.data(), y.data(), N, C, H, W, OC, k, k, st, st, pd, pd, OH, OW);
    dcnn_cpu_conv2d_bwd(x.data(), w.data(), dy.data(), dw.data(), dx.data(), N, C, H, W, OC, k, k, st, st, pd, pd, OH, OW);
    double ey = 0, ex = 0, ew = 0;
    for (size_t i = 0; i < y.size(); ++i) ey = fmax(ey, fabs(y[i] - yr[i]));
    for (size_t i = 0; i < dx.size(); ++i) ex = fmax(ex, fabs(dx[i] - dxr[i]));
    for (size_t i = 0; i < dw.size(); ++i) ew = fmax(ew, fabs(dw[i] - dwr[i]) / (1.0 + fabs(dwr[i])));
    CHECK(ey < 1e-4 && ex < 1e-4 && ew < 1e-4, "conv %dx%dx%dx%d oc%d k%d s%d p%d: %g %g %g", N, C, H, W, OC, k, st, pd, ey,
          ex, ew);
  }
}

int main(int argc, char** argv) {
  test_layers();
  test_conv();
  test_gemm();
  test_elementwise();
  test_im2col();
  test_augment();
  test_codec();
  test_loader(argc > 1 ? argv[1] : ".");
  failures += run_tcp_selftest();
  if (failures) {
    fprintf(stderr, "%d check(s) failed\n", failures);
    return 1;
  }
  printf("host selftest: all checks passed\n");
  return 0;
}
