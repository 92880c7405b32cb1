// Native CPU convolution, NCHW fp32: blocked im2col + GEMM with the output written in place.
// The reference's CPU Conv2D (src/nn/layers_impl/cpu/conv2d_ops.cpp:19-67) materialises im2col for the whole batch,
// runs one GEMM into a [OC][N*OH*OW] matrix and transposes that back to NCHW.  Here the batch is cut into blocks of
// images whose column matrix (K x 256..4096 floats, about 4 MB) stays cache-resident: a block is expanded, multiplied and written to
// its NCHW destination before the next one is touched, so neither the full column matrix nor the transposed output
// ever goes through memory.  Forward, data gradient and weight gradient share the block walk; the weight gradient
// re-expands the block instead of keeping the forward's columns (cheaper than reading them back from DRAM).
// Blocks are distributed over OpenMP threads (weight-gradient partials are summed in thread order); with fewer blocks
// than threads the blocks run one after the other and the GEMM inside parallelises itself.
#include <omp.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "host_common.h"

extern "C" void dcnn_cpu_sgemm(int transA, int transB, int M, int N, int K, float alpha, const float* A, int lda,
                               const float* B, int ldb, float beta, float* C, int ldc);

namespace {

struct Geo {
  int N, C, H, W, OC, kh, kw, sh, sw, ph, pw, OH, OW;
  int K() const { return C * kh * kw; }
  int P() const { return OH * OW; }
};

struct Scratch {
  std::vector<float> col, out, dw;
  float* need(std::vector<float>& v, size_t n) {
    if (v.size() < n) v.resize(n);
    return v.data();
  }
};
thread_local Scratch tls;

int images_per_block(const Geo& g) {
  // columns per block: as many as keep the K x columns expansion around 4 MB (L2-resident), between 256 (enough width
  // for the GEMM's register tiles to amortise packing) and 4096
  const long long budget_floats = 1 << 20;
  const int cols = (int)std::max<long long>(256, std::min<long long>(4096, budget_floats / std::max(1, g.K())));
  return std::max(1, std::min(g.N, cols / std::max(1, g.P())));
}

// output columns ow with 0 <= ow*sw - pw + s < W form the range [first_valid, end_valid)
int first_valid(const Geo& g, int s) { return g.pw > s ? (g.pw - s + g.sw - 1) / g.sw : 0; }
int end_valid(const Geo& g, int s) {
  const int last = g.W - 1 + g.pw - s;   // largest admissible ow*sw
  return last < 0 ? 0 : std::min(g.OW, last / g.sw + 1);
}

// col[K][nb*P] for images n0 .. n0+nb-1 (column index = i*P + oh*OW + ow, the reference's order within a block)
void expand_block(const Geo& g, const float* x, int n0, int nb, float* col) {
  const int P = g.P(), ld = nb * P;
  for (int c = 0; c < g.C; ++c)
    for (int r = 0; r < g.kh; ++r)
      for (int s = 0; s < g.kw; ++s) {
        float* row = col + (size_t)((c * g.kh + r) * g.kw + s) * ld;
        // valid ow range for this tap: 0 <= ow*sw - pw + s < W
        const int lo = first_valid(g, s), hi = end_valid(g, s);
        for (int i = 0; i < nb; ++i) {
          const float* src = x + ((size_t)(n0 + i) * g.C + c) * g.H * g.W;
          float* dst = row + (size_t)i * P;
          for (int oh = 0; oh < g.OH; ++oh, dst += g.OW) {
            const int ih = oh * g.sh - g.ph + r;
            if (ih < 0 || ih >= g.H || lo >= hi) {
              memset(dst, 0, sizeof(float) * g.OW);
              continue;
            }
            const float* line = src + (size_t)ih * g.W - g.pw + s;
            for (int ow = 0; ow < lo; ++ow) dst[ow] = 0.f;
            if (g.sw == 1) memcpy(dst + lo, line + lo, sizeof(float) * (hi - lo));
            else
              for (int ow = lo; ow < hi; ++ow) dst[ow] = line[ow * g.sw];
            for (int ow = hi; ow < g.OW; ++ow) dst[ow] = 0.f;
          }
        }
      }
}

// dx planes of images n0 .. n0+nb-1 = fold(col): every plane belongs to exactly one block, so plain adds suffice
void fold_block(const Geo& g, const float* col, int n0, int nb, float* dx) {
  const int P = g.P(), ld = nb * P;
  for (int i = 0; i < nb; ++i)
    for (int c = 0; c < g.C; ++c) {
      float* dst = dx + ((size_t)(n0 + i) * g.C + c) * g.H * g.W;
      memset(dst, 0, sizeof(float) * g.H * g.W);
      for (int r = 0; r < g.kh; ++r)
        for (int s = 0; s < g.kw; ++s) {
          const float* row = col + (size_t)((c * g.kh + r) * g.kw + s) * ld + (size_t)i * P;
          const int lo = first_valid(g, s), hi = end_valid(g, s);
          for (int oh = 0; oh < g.OH; ++oh) {
            const int ih = oh * g.sh - g.ph + r;
            if (ih < 0 || ih >= g.H) continue;
            float* line = dst + (size_t)ih * g.W - g.pw + s;
            const float* src = row + (size_t)oh * g.OW;
            if (g.sw == 1) {
#pragma omp simd
              for (int ow = lo; ow < hi; ++ow) line[ow] += src[ow];
            } else {
              for (int ow = lo; ow < hi; ++ow) line[ow * g.sw] += src[ow];
            }
          }
        }
    }
}

// out[OC][nb*P] <- dy images (identity when nb == 1: the image's [OC][P] block is already that matrix)
const float* gather_dy(const Geo& g, const float* dy, int n0, int nb, Scratch& sc) {
  const int P = g.P();
  if (nb == 1) return dy + (size_t)n0 * g.OC * P;
  float* out = sc.need(sc.out, (size_t)g.OC * nb * P);
  for (int oc = 0; oc < g.OC; ++oc)
    for (int i = 0; i < nb; ++i)
      memcpy(out + ((size_t)oc * nb + i) * P, dy + ((size_t)(n0 + i) * g.OC + oc) * P, sizeof(float) * P);
  return out;
}

template <typename F>
void for_blocks(const Geo& g, int nb, F&& body) {
  const int blocks = (g.N + nb - 1) / nb;
  const int threads = omp_get_max_threads();
  if (threads > 1 && blocks >= threads && !omp_in_parallel()) {
#pragma omp parallel for schedule(static)
    for (int b = 0; b < blocks; ++b) body(b * nb, std::min(nb, g.N - b * nb));
  } else {
    for (int b = 0; b < blocks; ++b) body(b * nb, std::min(nb, g.N - b * nb));
  }
}

}  // namespace

// y[N][OC][OH][OW] = conv(x[N][C][H][W], w[OC][C][kh][kw]) (+ bias[OC])
DCNN_HOST_API void dcnn_cpu_conv2d_fwd(const float* x, const float* w, const float* bias, float* y, int N, int C, int H,
                                       int W, int OC, int kh, int kw, int sh, int sw, int ph, int pw, int OH, int OW) {
  const Geo g{N, C, H, W, OC, kh, kw, sh, sw, ph, pw, OH, OW};
  const int K = g.K(), P = g.P(), nb = images_per_block(g);
  for_blocks(g, nb, [&](int n0, int cnt) {
    Scratch& sc = tls;
    float* col = sc.need(sc.col, (size_t)K * cnt * P);
    expand_block(g, x, n0, cnt, col);
    if (cnt == 1) {
      float* dst = y + (size_t)n0 * OC * P;
      dcnn_cpu_sgemm(0, 0, OC, P, K, 1.f, w, K, col, P, 0.f, dst, P);
      if (bias)
        for (int oc = 0; oc < OC; ++oc)
          for (int p = 0; p < P; ++p) dst[(size_t)oc * P + p] += bias[oc];
      return;
    }
    float* out = sc.need(sc.out, (size_t)OC * cnt * P);
    dcnn_cpu_sgemm(0, 0, OC, cnt * P, K, 1.f, w, K, col, cnt * P, 0.f, out, cnt * P);
    for (int i = 0; i < cnt; ++i)
      for (int oc = 0; oc < OC; ++oc) {
        float* dst = y + ((size_t)(n0 + i) * OC + oc) * P;
        const float* src = out + ((size_t)oc * cnt + i) * P;
        const float b = bias ? bias[oc] : 0.f;
        for (int p = 0; p < P; ++p) dst[p] = src[p] + b;
      }
  });
}

// dw[OC][C*kh*kw] += sum_n dy[n] * im2col(x[n])^T  (if dw)      dx = fold(w^T * dy)  (if dx; overwritten)
DCNN_HOST_API void dcnn_cpu_conv2d_bwd(const float* x, const float* w, const float* dy, float* dw, float* dx, int N, int C,
                                       int H, int W, int OC, int kh, int kw, int sh, int sw, int ph, int pw, int OH,
                                       int OW) {
  const Geo g{N, C, H, W, OC, kh, kw, sh, sw, ph, pw, OH, OW};
  const int K = g.K(), P = g.P(), nb = images_per_block(g);
  const int blocks = (N + nb - 1) / nb;
  const int threads = omp_get_max_threads();
  const bool par = threads > 1 && blocks >= threads && !omp_in_parallel();
  std::vector<float> partial;                       // per-thread weight-gradient partials when blocks run in parallel
  if (dw && par) partial.assign((size_t)threads * OC * K, 0.f);
  for_blocks(g, nb, [&](int n0, int cnt) {
    Scratch& sc = tls;
    float* col = sc.need(sc.col, (size_t)K * cnt * P);
    const float* gy = gather_dy(g, dy, n0, cnt, sc);
    if (dw) {
      expand_block(g, x, n0, cnt, col);
      float* acc = par ? partial.data() + (size_t)omp_get_thread_num() * OC * K : dw;
      dcnn_cpu_sgemm(0, 1, OC, K, cnt * P, 1.f, gy, cnt * P, col, cnt * P, 1.f, acc, K);
    }
    if (dx) {
      dcnn_cpu_sgemm(1, 0, K, cnt * P, OC, 1.f, w, K, gy, cnt * P, 0.f, col, cnt * P);
      fold_block(g, col, n0, cnt, dx);
    }
  });
  if (dw && par)
    for (int t = 0; t < threads; ++t) {
      const float* p = partial.data() + (size_t)t * OC * K;
      for (int i = 0; i < OC * K; ++i) dw[i] += p[i];
    }
}
