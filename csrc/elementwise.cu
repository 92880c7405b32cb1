// Pooling, activations, dropout, layout transposes and small elementwise kernels (NHWC fp32).
// Replaces the reference's maxpool_ops.cu / avgpool_ops.cu / dropout.cu / activations_impl/cuda/*.cu /
// ops/cuda/kernels.cu.
#include "common.cuh"

namespace dcnn {

// ---------------------------------------------------------------------------------------------- pooling
// One thread per (output pixel, 4 channels).  Stores the winning window slot (r*kw+s) as one byte instead of the
// reference's 8-byte flat index (maxpool_ops.cu:15-59); ties keep the first max in scan order.
// KH/KW > 0: compile-time window (3x3 and 2x2 cover the zoo), fully unrolled; 0: run-time window.  Index arithmetic is
// 32-bit (the host checks that the element counts fit): the 64-bit divisions of the first version cost more than the
// memory traffic (142 us for the ResNet-50 stem pool at batch 256 against a 54 us traffic floor).
template <int KH, int KW>
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                          uint8_t* __restrict__ idx, int N, int H, int W, int C, int OH,
                                                          int OW, int kh_, int kw_, int sh, int sw, int ph, int pw) {
  const int kh = KH > 0 ? KH : kh_, kw = KW > 0 ? KW : kw_;
  const unsigned C4 = (unsigned)C >> 2;
  const unsigned total = (unsigned)N * OH * OW * C4;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned c = (i % C4) * 4;
    unsigned t = i / C4;
    const int ow = (int)(t % (unsigned)OW); t /= (unsigned)OW;
    const int oh = (int)(t % (unsigned)OH);
    const unsigned n = t / (unsigned)OH;
    float4 best = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    uchar4 bi = make_uchar4(0, 0, 0, 0);
    const float* xn = x + (size_t)n * H * W * C + c;
#pragma unroll
    for (int r = 0; r < kh; ++r) {
      const int ih = oh * sh - ph + r;
      if ((unsigned)ih >= (unsigned)H) continue;
#pragma unroll
      for (int s = 0; s < kw; ++s) {
        const int iw = ow * sw - pw + s;
        if ((unsigned)iw >= (unsigned)W) continue;
        const float4 v = ldg4(xn + ((size_t)ih * W + iw) * C);
        const uint8_t slot = (uint8_t)(r * kw + s);
        if (v.x > best.x) { best.x = v.x; bi.x = slot; }
        if (v.y > best.y) { best.y = v.y; bi.y = slot; }
        if (v.z > best.z) { best.z = v.z; bi.z = slot; }
        if (v.w > best.w) { best.w = v.w; bi.w = slot; }
      }
    }
    st4(y + (size_t)i * 4, best);
    *reinterpret_cast<uchar4*>(idx + (size_t)i * 4) = bi;
  }
}

// Gather-style backward: one thread per (input pixel, 4 channels) sums the output positions that selected it.
// No atomics (the reference scatters with atomicAdd, maxpool_ops.cu:61-82) and no memset of dx.
template <int KH, int KW>
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ idx,
                                                          float* __restrict__ dx, int N, int H, int W, int C, int OH,
                                                          int OW, int kh_, int kw_, int sh, int sw, int ph, int pw) {
  const int kh = KH > 0 ? KH : kh_, kw = KW > 0 ? KW : kw_;
  const unsigned C4 = (unsigned)C >> 2;
  const unsigned total = (unsigned)N * H * W * C4;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned c = (i % C4) * 4;
    unsigned t = i / C4;
    const int w = (int)(t % (unsigned)W); t /= (unsigned)W;
    const int h = (int)(t % (unsigned)H);
    const unsigned n = t / (unsigned)H;
    float4 g = make_float4(0, 0, 0, 0);
    // output rows whose window covers h: oh*sh - ph <= h < oh*sh - ph + kh
    const int oh_lo = max(0, (h + ph - kh + sh) / sh), oh_hi = min(OH - 1, (h + ph) / sh);
    const int ow_lo = max(0, (w + pw - kw + sw) / sw), ow_hi = min(OW - 1, (w + pw) / sw);
    const size_t on = (size_t)n * OH * OW;
    for (int oh = oh_lo; oh <= oh_hi; ++oh) {
      const int r = h - (oh * sh - ph);
      if (r < 0 || r >= kh) continue;
      for (int ow = ow_lo; ow <= ow_hi; ++ow) {
        const int s = w - (ow * sw - pw);
        if (s < 0 || s >= kw) continue;
        const size_t o = (on + (size_t)oh * OW + ow) * C + c;
        const uchar4 bi = *reinterpret_cast<const uchar4*>(idx + o);
        const float4 d = ldg4(dy + o);
        const uint8_t slot = (uint8_t)(r * kw + s);
        if (bi.x == slot) g.x += d.x;
        if (bi.y == slot) g.y += d.y;
        if (bi.z == slot) g.z += d.z;
        if (bi.w == slot) g.w += d.w;
      }
    }
    st4(dx + (size_t)i * 4, g);
  }
}

// Backward of the 3x3 / stride 2 / pad 1 max pooling (the ResNet stem) by 2x2 INPUT blocks: the inputs (2a..2a+1,
// 2b..2b+1) can only have been selected by the four outputs (a..a+1, b..b+1), so one thread loads those four
// (index, gradient) pairs once -- 20 bytes per input float4 instead of 45 in the per-input gather -- and writes all four
// inputs.  Window slot of input (h, w) in output (oh, ow): (h - 2*oh + 1) * 3 + (w - 2*ow + 1).
__global__ void __launch_bounds__(256) maxpool_bwd_3x3s2p1_kernel(const float* __restrict__ dy,
                                                                  const uint8_t* __restrict__ idx,
                                                                  float* __restrict__ dx, int N, int H, int W, int C,
                                                                  int OH, int OW) {
  const unsigned C4 = (unsigned)C >> 2;
  const unsigned BH = (unsigned)(H + 1) >> 1, BW = (unsigned)(W + 1) >> 1;
  const unsigned total = (unsigned)N * BH * BW * C4;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned c = (i % C4) * 4;
    unsigned t = i / C4;
    const int b = (int)(t % BW); t /= BW;
    const int a = (int)(t % BH);
    const unsigned n = t / BH;
    float4 g[2][2];
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int v = 0; v < 2; ++v) g[u][v] = make_float4(0, 0, 0, 0);
#pragma unroll
    for (int p = 0; p < 2; ++p) {
      const int oh = a + p;
      if (oh >= OH) continue;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int ow = b + q;
        if (ow >= OW) continue;
        const size_t o = (((size_t)n * OH + oh) * OW + ow) * C + c;
        const uchar4 bi = *reinterpret_cast<const uchar4*>(idx + o);
        const float4 d = ldg4(dy + o);
        // inputs of this block inside output (oh, ow)'s window: rows 2a+u with u >= p (p = 1 only covers row 2a+1 = 2*oh-1),
        // likewise for the columns
#pragma unroll
        for (int u = p; u < 2; ++u)
#pragma unroll
          for (int v = q; v < 2; ++v) {
            const uint8_t slot = (uint8_t)((2 * a + u - 2 * oh + 1) * 3 + (2 * b + v - 2 * ow + 1));
            if (bi.x == slot) g[u][v].x += d.x;
            if (bi.y == slot) g[u][v].y += d.y;
            if (bi.z == slot) g[u][v].z += d.z;
            if (bi.w == slot) g[u][v].w += d.w;
          }
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int h = 2 * a + u;
      if (h >= H) continue;
#pragma unroll
      for (int v = 0; v < 2; ++v) {
        const int w = 2 * b + v;
        if (w >= W) continue;
        st4(dx + (((size_t)n * H + h) * W + w) * C + c, g[u][v]);
      }
    }
  }
}

// Average pooling; divisor always kh*kw, zero padding included (ref avgpool_ops.cu:106,122).
__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int N,
                                                          int H, int W, int C, int OH, int OW, int kh, int kw, int sh,
                                                          int sw, int ph, int pw) {
  const int C4 = C >> 2;
  const long long total = (long long)N * OH * OW * C4;
  const float inv = 1.f / (float)(kh * kw);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C4) * 4;
    long long t = i / C4;
    const int ow = (int)(t % OW); t /= OW;
    const int oh = (int)(t % OH);
    const int n = (int)(t / OH);
    float4 a = make_float4(0, 0, 0, 0);
    for (int r = 0; r < kh; ++r) {
      const int ih = oh * sh - ph + r;
      if ((unsigned)ih >= (unsigned)H) continue;
      for (int s = 0; s < kw; ++s) {
        const int iw = ow * sw - pw + s;
        if ((unsigned)iw >= (unsigned)W) continue;
        const float4 v = ldg4(x + (((long long)n * H + ih) * W + iw) * C + c);
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
      }
    }
    st4(y + i * 4, make_float4(a.x * inv, a.y * inv, a.z * inv, a.w * inv));
  }
}

__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, int N,
                                                          int H, int W, int C, int OH, int OW, int kh, int kw, int sh,
                                                          int sw, int ph, int pw) {
  const int C4 = C >> 2;
  const long long total = (long long)N * H * W * C4;
  const float inv = 1.f / (float)(kh * kw);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C4) * 4;
    long long t = i / C4;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H);
    const int n = (int)(t / H);
    float4 g = make_float4(0, 0, 0, 0);
    const int oh_lo = max(0, (h + ph - kh + sh) / sh), oh_hi = min(OH - 1, (h + ph) / sh);
    const int ow_lo = max(0, (w + pw - kw + sw) / sw), ow_hi = min(OW - 1, (w + pw) / sw);
    for (int oh = oh_lo; oh <= oh_hi; ++oh) {
      const int r = h - (oh * sh - ph);
      if (r < 0 || r >= kh) continue;
      for (int ow = ow_lo; ow <= ow_hi; ++ow) {
        const int s = w - (ow * sw - pw);
        if (s < 0 || s >= kw) continue;
        const float4 d = ldg4(dy + (((long long)n * OH + oh) * OW + ow) * C + c);
        g.x += d.x; g.y += d.y; g.z += d.z; g.w += d.w;
      }
    }
    st4(dx + i * 4, make_float4(g.x * inv, g.y * inv, g.z * inv, g.w * inv));
  }
}

// Scalar-channel pooling fallbacks (C % 4 != 0): mode 0 = max fwd, 1 = max bwd, 2 = avg fwd, 3 = avg bwd
__global__ void pool_scalar_kernel(int mode, const float* __restrict__ in, float* __restrict__ out,
                                   uint8_t* __restrict__ idx, int N, int H, int W, int C, int OH, int OW, int kh,
                                   int kw, int sh, int sw, int ph, int pw) {
  const bool fwd = (mode == 0 || mode == 2);
  const long long total = fwd ? (long long)N * OH * OW * C : (long long)N * H * W * C;
  const float inv = 1.f / (float)(kh * kw);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    long long t = i / C;
    if (fwd) {
      const int ow = (int)(t % OW); t /= OW;
      const int oh = (int)(t % OH);
      const int n = (int)(t / OH);
      float best = -INFINITY, acc = 0.f;
      uint8_t bi = 0;
      for (int r = 0; r < kh; ++r) {
        const int ih = oh * sh - ph + r;
        if ((unsigned)ih >= (unsigned)H) continue;
        for (int s = 0; s < kw; ++s) {
          const int iw = ow * sw - pw + s;
          if ((unsigned)iw >= (unsigned)W) continue;
          const float v = in[(((long long)n * H + ih) * W + iw) * C + c];
          acc += v;
          if (v > best) { best = v; bi = (uint8_t)(r * kw + s); }
        }
      }
      if (mode == 0) { out[i] = best; idx[i] = bi; } else { out[i] = acc * inv; }
    } else {
      const int w = (int)(t % W); t /= W;
      const int h = (int)(t % H);
      const int n = (int)(t / H);
      float g = 0.f;
      for (int oh = max(0, (h + ph - kh + sh) / sh); oh <= min(OH - 1, (h + ph) / sh); ++oh) {
        const int r = h - (oh * sh - ph);
        if (r < 0 || r >= kh) continue;
        for (int ow = max(0, (w + pw - kw + sw) / sw); ow <= min(OW - 1, (w + pw) / sw); ++ow) {
          const int s = w - (ow * sw - pw);
          if (s < 0 || s >= kw) continue;
          const long long o = (((long long)n * OH + oh) * OW + ow) * C + c;
          if (mode == 1) { if (idx[o] == (uint8_t)(r * kw + s)) g += in[o]; } else { g += in[o] * inv; }
        }
      }
      out[i] = g;
    }
  }
}

// ---------------------------------------------------------------------------------------------- activations
enum ActKind { ACT_RELU = 0, ACT_LEAKY = 1, ACT_ELU = 2, ACT_SIGMOID = 3, ACT_TANH = 4, ACT_LINEAR = 5 };

template <int KIND>
__device__ __forceinline__ float act_fwd(float x, float a) {
  if (KIND == ACT_RELU) return fmaxf(x, 0.f);
  if (KIND == ACT_LEAKY) return x > 0.f ? x : a * x;
  if (KIND == ACT_ELU) return x > 0.f ? x : a * (__expf(x) - 1.f);
  if (KIND == ACT_SIGMOID) return 1.f / (1.f + __expf(-x));
  if (KIND == ACT_TANH) return tanhf(x);
  return x;
}
template <int KIND>
__device__ __forceinline__ float act_bwd(float x, float g, float a) {  // x = pre-activation input
  if (KIND == ACT_RELU) return x > 0.f ? g : 0.f;
  if (KIND == ACT_LEAKY) return x > 0.f ? g : a * g;
  if (KIND == ACT_ELU) return x > 0.f ? g : g * a * __expf(x);
  if (KIND == ACT_SIGMOID) { const float s = 1.f / (1.f + __expf(-x)); return g * s * (1.f - s); }
  if (KIND == ACT_TANH) { const float t = tanhf(x); return g * (1.f - t * t); }
  return g;
}

template <int KIND, bool BWD>
__global__ void __launch_bounds__(256) act_kernel(const float* __restrict__ x, const float* __restrict__ g,
                                                  float* __restrict__ y, long long n, float a) {
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = ldcs4(x + i * 4);
    float4 o;
    if (BWD) {
      const float4 d = ldcs4(g + i * 4);
      o.x = act_bwd<KIND>(v.x, d.x, a); o.y = act_bwd<KIND>(v.y, d.y, a);
      o.z = act_bwd<KIND>(v.z, d.z, a); o.w = act_bwd<KIND>(v.w, d.w, a);
    } else {
      o.x = act_fwd<KIND>(v.x, a); o.y = act_fwd<KIND>(v.y, a); o.z = act_fwd<KIND>(v.z, a); o.w = act_fwd<KIND>(v.w, a);
    }
    st4(y + i * 4, o);
  }
  // tail
  const long long base = n4 << 2;
  const long long t = base + blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t < n) y[t] = BWD ? act_bwd<KIND>(x[t], g[t], a) : act_fwd<KIND>(x[t], a);
}

// Channel softmax over rows of length C (NHWC: one row per pixel); one warp per row.
__global__ void softmax_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long long rows, int C) {
  const long long row = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const float* xr = x + row * C;
  float m = -INFINITY;
  for (int c = lane; c < C; c += 32) m = fmaxf(m, xr[c]);
  m = warp_max(m);
  float s = 0.f;
  for (int c = lane; c < C; c += 32) s += __expf(xr[c] - m);
  s = warp_sum(s);
  const float inv = 1.f / s;
  for (int c = lane; c < C; c += 32) y[row * C + c] = __expf(xr[c] - m) * inv;
}
__global__ void softmax_bwd_kernel(const float* __restrict__ x, const float* __restrict__ g, float* __restrict__ dx,
                                   long long rows, int C) {
  const long long row = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const float* xr = x + row * C;
  const float* gr = g + row * C;
  float m = -INFINITY;
  for (int c = lane; c < C; c += 32) m = fmaxf(m, xr[c]);
  m = warp_max(m);
  float s = 0.f, dot = 0.f;
  for (int c = lane; c < C; c += 32) {
    const float e = __expf(xr[c] - m);
    s += e; dot = fmaf(e, gr[c], dot);
  }
  s = warp_sum(s); dot = warp_sum(dot);
  const float inv = 1.f / s;
  dot *= inv;
  for (int c = lane; c < C; c += 32) dx[row * C + c] = __expf(xr[c] - m) * inv * (gr[c] - dot);
}

__global__ void __launch_bounds__(256) add_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                  float* __restrict__ y, long long n) {
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 u = ldcs4(a + i * 4), v = ldcs4(b + i * 4);
    st4(y + i * 4, make_float4(u.x + v.x, u.y + v.y, u.z + v.z, u.w + v.w));
  }
  const long long t = (n4 << 2) + blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t < n) y[t] = a[t] + b[t];
}

// ---------------------------------------------------------------------------------------------- dropout
// Counter-based RNG (Philox4x32-10): element i of call `offset` always sees the same random number, so backward
// regenerates the mask instead of storing it (the reference stores a full mask tensor, dropout.cu:28-125).
__device__ __forceinline__ uint4 philox4x32(uint4 ctr, uint2 key) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += 0x9E3779B9u; key.y += 0xBB67AE85u;
  }
  return ctr;
}
__global__ void __launch_bounds__(256) dropout_kernel(const float* __restrict__ x, float* __restrict__ y, long long n,
                                                      float rate, unsigned long long seed, unsigned long long offset) {
  const float scale = 1.f / (1.f - rate);
  const long long n4 = (n + 3) >> 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const uint4 r = philox4x32(make_uint4((uint32_t)i, (uint32_t)(i >> 32), (uint32_t)offset, (uint32_t)(offset >> 32)),
                               make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long e = i * 4 + j;
      if (e < n) {
        const float u = (float)(rr[j] >> 8) * (1.f / 16777216.f);
        y[e] = u >= rate ? x[e] * scale : 0.f;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------- layout
// NHWC <-> NCHW per image: in [rows][cols] -> out [cols][rows], batched over blockIdx.z.
__global__ void transpose_batched_kernel(const float* __restrict__ in, float* __restrict__ out, int rows, int cols) {
  __shared__ float tile[32][33];
  const long long off = (long long)blockIdx.z * rows * cols;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int r = r0 + j, c = c0 + threadIdx.x;
    if (r < rows && c < cols) tile[j][threadIdx.x] = in[off + (long long)r * cols + c];
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int c = c0 + j, r = r0 + threadIdx.x;
    if (r < rows && c < cols) out[off + (long long)c * rows + r] = tile[threadIdx.x][j];
  }
}

}  // namespace dcnn
using namespace dcnn;

static int pool_launch(int mode, const float* in, float* out, uint8_t* idx, int N, int H, int W, int C, int OH, int OW,
                       int kh, int kw, int sh, int sw, int ph, int pw, cudaStream_t stream) {
  DCNN_REQUIRE(kh * kw <= 255, "pool window too large");
  const bool fwd = (mode == 0 || mode == 2);
  const long long total = (fwd ? (long long)N * OH * OW : (long long)N * H * W) * C;
  if ((C & 3) == 0) {
    const int grid = ew_grid(total / 4, 256);
    if (mode <= 1)
      DCNN_REQUIRE((long long)N * H * W * (C / 4) < (1ll << 32) && (long long)N * OH * OW * (C / 4) < (1ll << 32),
                   "max pooling: more than 2^32 float4 elements");
    if (mode == 0) {
      if (kh == 3 && kw == 3) maxpool_fwd_kernel<3, 3><<<grid, 256, 0, stream>>>(in, out, idx, N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw);
      else if (kh == 2 && kw == 2) maxpool_fwd_kernel<2, 2><<<grid, 256, 0, stream>>>(in, out, idx, N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw);
      else maxpool_fwd_kernel<0, 0><<<grid, 256, 0, stream>>>(in, out, idx, N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw);
    } else if (mode == 1) {
      if (kh == 3 && kw == 3 && sh == 2 && sw == 2 && ph == 1 && pw == 1 && OH == (H - 1) / 2 + 1 && OW == (W - 1) / 2 + 1) {
        const long long blocks4 = (long long)N * ((H + 1) / 2) * ((W + 1) / 2) * (C / 4);
        maxpool_bwd_3x3s2p1_kernel<<<ew_grid(blocks4, 256), 256, 0, stream>>>(in, idx, out, N, H, W, C, OH, OW);
      } else if (kh == 3 && kw == 3) maxpool_bwd_kernel<3, 3><<<grid, 256, 0, stream>>>(in, idx, out, N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw);
      else if (kh == 2 && kw == 2) maxpool_bwd_kernel<2, 2><<<grid, 256, 0, stream>>>(in, idx, out, N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw);
      else maxpool_bwd_kernel<0, 0><<<grid, 256, 0, stream>>>(in, idx, out, N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw);
    }
    else if (mode == 2) avgpool_fwd_kernel<<<grid, 256, 0, stream>>>(in, out, N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw);
    else avgpool_bwd_kernel<<<grid, 256, 0, stream>>>(in, out, N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw);
  } else {
    pool_scalar_kernel<<<ew_grid(total, 256), 256, 0, stream>>>(mode, in, out, idx, N, H, W, C, OH, OW, kh, kw, sh, sw,
                                                                ph, pw);
  }
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_maxpool_fwd(const float* x, float* y, uint8_t* idx, int N, int H, int W, int C, int OH, int OW,
                              int kh, int kw, int sh, int sw, int ph, int pw, cudaStream_t stream) {
  return pool_launch(0, x, y, idx, N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw, stream);
}
DCNN_API int dcnn_maxpool_bwd(const float* dy, const uint8_t* idx, float* dx, int N, int H, int W, int C, int OH,
                              int OW, int kh, int kw, int sh, int sw, int ph, int pw, cudaStream_t stream) {
  return pool_launch(1, dy, dx, const_cast<uint8_t*>(idx), N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw, stream);
}
DCNN_API int dcnn_avgpool_fwd(const float* x, float* y, int N, int H, int W, int C, int OH, int OW, int kh, int kw,
                              int sh, int sw, int ph, int pw, cudaStream_t stream) {
  return pool_launch(2, x, y, nullptr, N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw, stream);
}
DCNN_API int dcnn_avgpool_bwd(const float* dy, float* dx, int N, int H, int W, int C, int OH, int OW, int kh, int kw,
                              int sh, int sw, int ph, int pw, cudaStream_t stream) {
  return pool_launch(3, dy, dx, nullptr, N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw, stream);
}

template <bool BWD>
static int act_launch(int kind, const float* x, const float* g, float* y, long long n, float alpha,
                      cudaStream_t stream) {
  const int grid = ew_grid((n >> 2) + 1, 256);
  switch (kind) {
    case ACT_RELU: act_kernel<ACT_RELU, BWD><<<grid, 256, 0, stream>>>(x, g, y, n, alpha); break;
    case ACT_LEAKY: act_kernel<ACT_LEAKY, BWD><<<grid, 256, 0, stream>>>(x, g, y, n, alpha); break;
    case ACT_ELU: act_kernel<ACT_ELU, BWD><<<grid, 256, 0, stream>>>(x, g, y, n, alpha); break;
    case ACT_SIGMOID: act_kernel<ACT_SIGMOID, BWD><<<grid, 256, 0, stream>>>(x, g, y, n, alpha); break;
    case ACT_TANH: act_kernel<ACT_TANH, BWD><<<grid, 256, 0, stream>>>(x, g, y, n, alpha); break;
    case ACT_LINEAR: act_kernel<ACT_LINEAR, BWD><<<grid, 256, 0, stream>>>(x, g, y, n, alpha); break;
    default: DCNN_REQUIRE(false, "unknown activation kind %d", kind);
  }
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}
DCNN_API int dcnn_act_fwd(int kind, const float* x, float* y, long long n, float alpha, cudaStream_t stream) {
  return act_launch<false>(kind, x, nullptr, y, n, alpha, stream);
}
DCNN_API int dcnn_act_bwd(int kind, const float* x, const float* g, float* dx, long long n, float alpha,
                          cudaStream_t stream) {
  return act_launch<true>(kind, x, g, dx, n, alpha, stream);
}
DCNN_API int dcnn_softmax_fwd(const float* x, float* y, long long rows, int C, cudaStream_t stream) {
  softmax_fwd_kernel<<<ceil_div(rows, 8), 256, 0, stream>>>(x, y, rows, C);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}
DCNN_API int dcnn_softmax_bwd(const float* x, const float* g, float* dx, long long rows, int C, cudaStream_t stream) {
  softmax_bwd_kernel<<<ceil_div(rows, 8), 256, 0, stream>>>(x, g, dx, rows, C);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}
DCNN_API int dcnn_add(const float* a, const float* b, float* y, long long n, cudaStream_t stream) {
  add_kernel<<<ew_grid((n >> 2) + 1, 256), 256, 0, stream>>>(a, b, y, n);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}
DCNN_API int dcnn_dropout(const float* x, float* y, long long n, float rate, unsigned long long seed,
                          unsigned long long offset, cudaStream_t stream) {
  dropout_kernel<<<ew_grid((n + 3) >> 2, 256), 256, 0, stream>>>(x, y, n, rate, seed, offset);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}
// out[b][c][r] = in[b][r][c]
DCNN_API int dcnn_transpose_batched(const float* in, float* out, int batch, int rows, int cols, cudaStream_t stream) {
  dim3 grid(ceil_div(cols, 32), ceil_div(rows, 32), batch), block(32, 8);
  transpose_batched_kernel<<<grid, block, 0, stream>>>(in, out, rows, cols);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}
