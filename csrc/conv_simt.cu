// CUDA-core (SIMT, exact FP32) implicit-GEMM fallbacks for shapes the tensor-core path does not take:
// channel counts that are not a multiple of 32 (the RGB stem, the MNIST CNN, odd Dense widths).
// One tiled gather-GEMM kernel, three instantiations: fprop, dgrad, wgrad (split-K with FP32 atomics so that
// weight gradients accumulate, beta=1, like the reference's cudnnConvolutionBackwardFilter call).
// Also: the weight re-layout used by the tensor-core dgrad, and the exact-FP32 Dense paths.
#include "common.cuh"

namespace dcnn {

struct ConvGeom {
  int N, H, W, C;     // input
  int OH, OW, K;      // output (K = out channels)
  int R, S, sh, sw, ph, pw;
};

// ---- operand functors: element (row, k) of the logical GEMM operand, 0 when masked -------------------
struct FpropA {  // row = output pixel, k = (r,s,c)
  const float* x; ConvGeom g;
  __device__ __forceinline__ float operator()(int m, int k) const {
    const int c = k % g.C; const int t = k / g.C; const int s = t % g.S; const int r = t / g.S;
    const int ow = m % g.OW; const int oh = (m / g.OW) % g.OH; const int n = m / (g.OW * g.OH);
    const int ih = oh * g.sh - g.ph + r, iw = ow * g.sw - g.pw + s;
    if ((unsigned)ih >= (unsigned)g.H || (unsigned)iw >= (unsigned)g.W) return 0.f;
    return __ldg(x + (((long long)n * g.H + ih) * g.W + iw) * g.C + c);
  }
};
struct FpropB {  // row = oc, k = (r,s,c): weights are [K][R][S][C]
  const float* w; int Ktot;
  __device__ __forceinline__ float operator()(int n, int k) const { return __ldg(w + (long long)n * Ktot + k); }
};
struct DgradA {  // row = input pixel (n,h,w), k = (r,s,oc)
  const float* dy; ConvGeom g;
  __device__ __forceinline__ float operator()(int m, int k) const {
    const int oc = k % g.K; const int t = k / g.K; const int s = t % g.S; const int r = t / g.S;
    const int w_ = m % g.W; const int h = (m / g.W) % g.H; const int n = m / (g.W * g.H);
    const int a = h + g.ph - r, b = w_ + g.pw - s;
    if (a < 0 || b < 0 || a % g.sh || b % g.sw) return 0.f;
    const int oh = a / g.sh, ow = b / g.sw;
    if (oh >= g.OH || ow >= g.OW) return 0.f;
    return __ldg(dy + (((long long)n * g.OH + oh) * g.OW + ow) * g.K + oc);
  }
};
struct DgradB {  // row = c, k = (r,s,oc)
  const float* w; ConvGeom g;
  __device__ __forceinline__ float operator()(int c, int k) const {
    const int oc = k % g.K; const int t = k / g.K;
    return __ldg(w + ((long long)oc * g.R * g.S + t) * g.C + c);
  }
};
struct WgradA {  // row = oc, k = output pixel
  const float* dy; int K;
  __device__ __forceinline__ float operator()(int oc, int pix) const { return __ldg(dy + (long long)pix * K + oc); }
};
struct WgradB {  // row = (r,s,c), k = output pixel
  const float* x; ConvGeom g;
  __device__ __forceinline__ float operator()(int j, int pix) const {
    const int c = j % g.C; const int t = j / g.C; const int s = t % g.S; const int r = t / g.S;
    const int ow = pix % g.OW; const int oh = (pix / g.OW) % g.OH; const int n = pix / (g.OW * g.OH);
    const int ih = oh * g.sh - g.ph + r, iw = ow * g.sw - g.pw + s;
    if ((unsigned)ih >= (unsigned)g.H || (unsigned)iw >= (unsigned)g.W) return 0.f;
    return __ldg(x + (((long long)n * g.H + ih) * g.W + iw) * g.C + c);
  }
};

// C[m][n] = sum_k A(m,k) * B(n,k); 64x64 tile, 16-deep k slices, 256 threads, 4x4 outputs per thread.
// ATOMIC: split-K over blockIdx.z, results added with atomics (accumulates into C).
template <class FA, class FB, bool ATOMIC>
__global__ void __launch_bounds__(256) gather_gemm_kernel(int M, int N, int K, int k_per_split, FA fa, FB fb,
                                                          float* __restrict__ C, long long ldc,
                                                          const float* __restrict__ bias, int accumulate) {
  __shared__ float sA[16][64 + 4];
  __shared__ float sB[16][64 + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.x * 64, n0 = blockIdx.y * 64;
  const int kbeg = blockIdx.z * k_per_split;
  const int kend = min(K, kbeg + k_per_split);
  float acc[4][4] = {};
  for (int k0 = kbeg; k0 < kend; k0 += 16) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = tid + i * 256;  // 0..1023
      const int kk = idx & 15, row = idx >> 4;
      const int k = k0 + kk;
      sA[kk][row] = (m0 + row < M && k < kend) ? fa(m0 + row, k) : 0.f;
      sB[kk][row] = (n0 + row < N && k < kend) ? fb(n0 + row, k) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = sA[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = sB[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      float* dst = C + (long long)m * ldc + n;
      if (ATOMIC) {
        atomicAdd(dst, acc[i][j]);
      } else {
        float v = acc[i][j] + (bias ? __ldg(bias + n) : 0.f);
        if (accumulate) v += *dst;
        *dst = v;
      }
    }
  }
}

// W[K][R][S][C]  ->  Wt[C][R][S][K] with both spatial axes flipped: the weights of the stride-1 dgrad
// expressed as a forward convolution over dY.
__global__ void weight_dgrad_layout_kernel(const float* __restrict__ w, float* __restrict__ wt, int K, int R, int S,
                                           int C) {
  const long long total = (long long)K * R * S * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    // i indexes the destination [c][r'][s'][k]
    const int k = (int)(i % K);
    long long t = i / K;
    const int s2 = (int)(t % S); t /= S;
    const int r2 = (int)(t % R);
    const int c = (int)(t / R);
    wt[i] = __ldg(w + (((long long)k * R + (R - 1 - r2)) * S + (S - 1 - s2)) * C + c);
  }
}


// Up to kWtBatch weight tensors in ONE launch: the table {w, wt, K, R, S, C} travels as a kernel parameter (no device
// table, so the launch is capturable without a host-to-device copy); the CTAs stride over the 32 x 32 tiles of all entries.
constexpr int kWtBatch = 64;
struct WtTable {
  long long e[kWtBatch][6];
  int tile_end[kWtBatch];   // running total of 32 x 32 tiles up to and including entry i
  int count;
};
__global__ void __launch_bounds__(256) weight_dgrad_layout_batched_kernel(const __grid_constant__ WtTable table) {
  // per tap a [K][C] -> [C][K] transpose (taps mirrored) through a 32 x 32 shared-memory tile: reads coalesced along c,
  // writes coalesced along k, and the index arithmetic is per tile instead of four 64-bit divisions per element (the
  // element-wise version ran at 1.5 TB/s: 128 us for the 94 MB of ResNet-50 weights)
  // The tiles of all entries form one list that the CTAs stride over (an entry per blockIdx.y with a fixed number of
  // CTAs left the three 2.4M-element weights of the last stage to 64 CTAs each: 119 us, latency-bound).
  __shared__ float tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int total_tiles = table.tile_end[table.count - 1];
  for (int g = blockIdx.x; g < total_tiles; g += gridDim.x) {
    int lo = 0, hi = table.count - 1;   // first entry whose running total exceeds g
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (table.tile_end[mid] > g) hi = mid; else lo = mid + 1;
    }
    const long long* e = table.e[lo];
    const float* __restrict__ w = reinterpret_cast<const float*>(e[0]);
    float* __restrict__ wt = reinterpret_cast<float*>(e[1]);
    const int K = (int)e[2], R = (int)e[3], S = (int)e[4], C = (int)e[5];
    const int kt = (K + 31) >> 5, ct = (C + 31) >> 5;
    const int t = g - (lo > 0 ? table.tile_end[lo - 1] : 0);
    const int cb = t % ct;
    const int u = t / ct;
    const int kb = u % kt, tap = u / kt;
    const int r = tap / S, s2 = tap % S;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = kb * 32 + ty + 8 * j, c = cb * 32 + tx;
      tile[ty + 8 * j][tx] = (k < K && c < C) ? __ldg(w + (((long long)k * R + r) * S + s2) * C + c) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = cb * 32 + ty + 8 * j, k = kb * 32 + tx;
      if (c < C && k < K) wt[(((long long)c * R + (R - 1 - r)) * S + (S - 1 - s2)) * K + k] = tile[tx][ty + 8 * j];
    }
    __syncthreads();
  }
}

}  // namespace dcnn
using namespace dcnn;

static ConvGeom make_geom(int N, int H, int W, int C, int K, int R, int S, int sh, int sw, int ph, int pw) {
  ConvGeom g;
  g.N = N; g.H = H; g.W = W; g.C = C; g.K = K; g.R = R; g.S = S; g.sh = sh; g.sw = sw; g.ph = ph; g.pw = pw;
  g.OH = (H + 2 * ph - R) / sh + 1;
  g.OW = (W + 2 * pw - S) / sw + 1;
  return g;
}

DCNN_API int dcnn_conv_fprop_simt(const float* x, const float* w, const float* bias, float* y, int N, int H, int W,
                                  int C, int K, int R, int S, int sh, int sw, int ph, int pw, cudaStream_t stream) {
  ConvGeom g = make_geom(N, H, W, C, K, R, S, sh, sw, ph, pw);
  const long long M = (long long)N * g.OH * g.OW;
  const int Ktot = R * S * C;
  dim3 grid(ceil_div(M, 64), ceil_div(K, 64), 1);
  gather_gemm_kernel<FpropA, FpropB, false><<<grid, 256, 0, stream>>>((int)M, K, Ktot, Ktot, FpropA{x, g},
                                                                      FpropB{w, Ktot}, y, K, bias, 0);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_conv_dgrad_simt(const float* dy, const float* w, float* dx, int N, int H, int W, int C, int K, int R,
                                  int S, int sh, int sw, int ph, int pw, int accumulate, cudaStream_t stream) {
  ConvGeom g = make_geom(N, H, W, C, K, R, S, sh, sw, ph, pw);
  const long long M = (long long)N * H * W;
  const int Kred = R * S * K;
  dim3 grid(ceil_div(M, 64), ceil_div(C, 64), 1);
  gather_gemm_kernel<DgradA, DgradB, false><<<grid, 256, 0, stream>>>((int)M, C, Kred, Kred, DgradA{dy, g},
                                                                      DgradB{w, g}, dx, C, nullptr, accumulate);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_conv_wgrad_simt(const float* dy, const float* x, float* dw, int N, int H, int W, int C, int K, int R,
                                  int S, int sh, int sw, int ph, int pw, cudaStream_t stream) {
  ConvGeom g = make_geom(N, H, W, C, K, R, S, sh, sw, ph, pw);
  const long long P = (long long)N * g.OH * g.OW;
  const int Ktot = R * S * C;
  const int tiles = ceil_div(K, 64) * ceil_div(Ktot, 64);
  int splits = ceil_div(4 * num_sms(), tiles);
  const int max_splits = ceil_div(P, 256);
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  int per = ceil_div(P, splits);
  per = ceil_div(per, 16) * 16;
  splits = ceil_div(P, per);
  dim3 grid(ceil_div(K, 64), ceil_div(Ktot, 64), splits);
  gather_gemm_kernel<WgradA, WgradB, true><<<grid, 256, 0, stream>>>(K, Ktot, (int)P, per, WgradA{dy, K},
                                                                     WgradB{x, g}, dw, Ktot, nullptr, 1);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_weight_dgrad_layout(const float* w, float* wt, int K, int R, int S, int C, cudaStream_t stream) {
  const long long total = (long long)K * R * S * C;
  weight_dgrad_layout_kernel<<<ew_grid(total, 256), 256, 0, stream>>>(w, wt, K, R, S, C);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_weight_dgrad_layout_batched(const long long* table_host, int count, cudaStream_t stream) {
  DCNN_REQUIRE(count >= 1, "dcnn_weight_dgrad_layout_batched: bad entry count %d", count);
  for (int first = 0; first < count; first += kWtBatch) {
    const int n = count - first < kWtBatch ? count - first : kWtBatch;
    WtTable t;
    memset(&t, 0, sizeof(t));
    memcpy(t.e, table_host + 6 * (size_t)first, sizeof(long long) * 6 * n);
    long long total = 0;
    for (int i = 0; i < n; ++i) {
      const long long K = t.e[i][2], R = t.e[i][3], S = t.e[i][4], C = t.e[i][5];
      total += R * S * ((K + 31) / 32) * ((C + 31) / 32);
      DCNN_REQUIRE(total < (1ll << 31), "dcnn_weight_dgrad_layout_batched: too many tiles (%lld)", total);
      t.tile_end[i] = (int)total;
    }
    t.count = n;
    const long long cap = 8ll * num_sms();
    weight_dgrad_layout_batched_kernel<<<(unsigned)(total < cap ? total : cap), 256, 0, stream>>>(t);
    DCNN_CHECK_LAUNCH();
    DCNN_COUNT_LAUNCH(1);
  }
  return 0;
}

namespace dcnn {
// Packed-tap im2col for small-Cin ("stem") convolutions: P[n,oh,ow,k] = x[n, oh*sh-ph+r, ow*sw-pw+s, c] with
// k = (r*S+s)*Cin + c for k < R*S*Cin, zero for the padding channels up to Kp (a multiple of 32) and for taps that fall
// outside the image.  The stem then runs as a 1x1 convolution with Cin = Kp on the tensor-core kernels instead of on
// CUDA cores (a 3x3x3 stem is K = 27 -> 32: no wasted MMA work to speak of).  k is the fastest index: stores are
// full 128-byte rows; the reads hit L1/L2 (every input element is read R*S times).
__global__ void __launch_bounds__(256) pack_taps_kernel(const float* __restrict__ x, float* __restrict__ P, int N, int H,
                                                        int W, int Cin, int OH, int OW, int R, int S, int sh, int sw,
                                                        int ph, int pw, int Kp) {
  // tab[k] = (r << 24) | (s << 16) | c for the packed channel k, -1 for the zero padding up to Kp
  __shared__ int tab[256];
  const int K = R * S * Cin;
  for (int k = threadIdx.x; k < Kp; k += blockDim.x) {
    int v = -1;
    if (k < K) {
      const int tap = k / Cin, c = k - tap * Cin;
      const int r = tap / S, s2 = tap - r * S;
      v = (r << 24) | (s2 << 16) | c;
    }
    tab[k] = v;
  }
  __syncthreads();
  // a thread packs 4 consecutive k of one pixel (one 16-byte store); Q = Kp/4 threads cover a pixel, so consecutive
  // threads write whole 4*Kp-byte rows back to back.  Threads beyond the last whole pixel of the block idle.
  const int Q = Kp >> 2;
  const int q = threadIdx.x % Q;
  const int pix_per_block = blockDim.x / Q;
  const long long pixels = (long long)N * OH * OW;
  if ((int)threadIdx.x >= pix_per_block * Q) return;
  for (long long pix = (long long)blockIdx.x * pix_per_block + threadIdx.x / Q; pix < pixels;
       pix += (long long)gridDim.x * pix_per_block) {
    const unsigned up = (unsigned)pix;                       // N*OH*OW < 2^32 (checked on the host)
    const unsigned ow = up % (unsigned)OW, t = up / (unsigned)OW;
    const unsigned oh = t % (unsigned)OH, n = t / (unsigned)OH;
    const int ih0 = (int)oh * sh - ph, iw0 = (int)ow * sw - pw;
    const float* xin = x + (long long)n * H * W * Cin;
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int e = tab[4 * q + j];
      v[j] = 0.f;
      if (e >= 0) {
        const int ih = ih0 + (e >> 24), iw = iw0 + ((e >> 16) & 255);
        if ((unsigned)ih < (unsigned)H && (unsigned)iw < (unsigned)W) v[j] = __ldg(xin + ((long long)ih * W + iw) * Cin + (e & 0xffff));
      }
    }
    *reinterpret_cast<float4*>(P + pix * Kp + 4 * q) = make_float4(v[0], v[1], v[2], v[3]);
  }
}
}  // namespace dcnn

DCNN_API int dcnn_pack_taps(const float* x, float* P, int N, int H, int W, int Cin, int OH, int OW, int R, int S,
                            int sh, int sw, int ph, int pw, int Kp, cudaStream_t stream) {
  DCNN_REQUIRE(Kp % 32 == 0 && Kp >= R * S * Cin, "dcnn_pack_taps: Kp (%d) must be a multiple of 32 >= R*S*Cin (%d)", Kp,
               R * S * Cin);
  DCNN_REQUIRE(Kp <= 256, "dcnn_pack_taps: Kp (%d) must not exceed 256", Kp);
  DCNN_REQUIRE((long long)N * OH * OW < (1ll << 32) && R < 128 && S < 256 && Cin < 65536, "dcnn_pack_taps: problem too large");
  const long long pixels = (long long)N * OH * OW;
  const int ppb = 256 / (Kp / 4);
  long long blocks = (pixels + ppb - 1) / ppb;
  const long long cap = (long long)dcnn::num_sms() * 16;
  if (blocks > cap) blocks = cap;
  dcnn::pack_taps_kernel<<<(unsigned)blocks, 256, 0, stream>>>(x, P, N, H, W, Cin, OH, OW, R, S, sh, sw, ph, pw, Kp);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}
