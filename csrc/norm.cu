// BatchNorm / GroupNorm kernels for NHWC fp32 activations (x viewed as [M = N*H*W rows][C channels]).
//
// BatchNorm forward statistics are *partials*: per 128-row tile (sum, M2 about the tile mean) per channel plus a
// row count per tile.  The tensor-core conv epilogue (conv_tc.cu) emits exactly this format, so for conv->BN the
// statistics cost no extra pass over the activation; bn_colstats_kernel produces it for every other producer.
// bn_finalize merges partials with Chan's parallel-variance formula in FP64 (numerically equivalent to the
// reference's Welford kernel, batchnorm_ops.cu:108-155), updates the running statistics (unbiased variance,
// momentum) and precomputes scale/shift so the apply pass is one FMA per element with ReLU / residual fused.
// Backward = one reduce pass (ReLU mask folded in) + finalize + one apply pass (dx = g*k1 + x*k2 + k3).
#include <atomic>

#include "peer.cuh"

namespace dcnn {

constexpr int kStatRows = 128;  // rows per partial tile of the standalone statistics kernel

// grid: (row_tiles, ceil(C/128)); block (32, 8): lane -> 4 channels (float4), ty strides rows.
__global__ void __launch_bounds__(256) bn_colstats_kernel(const float* __restrict__ x, long long M, int C,
                                                          int rows_per_tile, int tiles, float* __restrict__ partials) {
  __shared__ float4 s_sum[8][32];
  __shared__ float4 s_sq[8][32];
  const int tile = blockIdx.x;
  const int c = (blockIdx.y * 32 + threadIdx.x) * 4;
  const long long r0 = (long long)tile * rows_per_tile;
  const long long r1 = min(M, r0 + rows_per_tile);
  const bool active = c < C;
  float4 shift = make_float4(0, 0, 0, 0), s = shift, q = shift;
  if (active) {
    shift = ldg4(x + r0 * C + c);  // shifted sums: robust against |mean| >> std
#pragma unroll 4
    for (long long r = r0 + threadIdx.y; r < r1; r += 8) {
      const float4 v = ldcs4(x + r * C + c);
      const float dx = v.x - shift.x, dy = v.y - shift.y, dz = v.z - shift.z, dw = v.w - shift.w;
      s.x += dx; s.y += dy; s.z += dz; s.w += dw;
      q.x = fmaf(dx, dx, q.x); q.y = fmaf(dy, dy, q.y); q.z = fmaf(dz, dz, q.z); q.w = fmaf(dw, dw, q.w);
    }
  }
  s_sum[threadIdx.y][threadIdx.x] = s;
  s_sq[threadIdx.y][threadIdx.x] = q;
  __syncthreads();
  if (threadIdx.y == 0 && active) {
    for (int i = 1; i < 8; ++i) {
      const float4 a = s_sum[i][threadIdx.x], b = s_sq[i][threadIdx.x];
      s.x += a.x; s.y += a.y; s.z += a.z; s.w += a.w;
      q.x += b.x; q.y += b.y; q.z += b.z; q.w += b.w;
    }
    const float n = (float)(r1 - r0);
    float* ps = partials + ((long long)tile * 2 + 0) * C + c;
    float* pm = partials + ((long long)tile * 2 + 1) * C + c;
    // sum = sum(d) + n*shift ; M2 = sum(d^2) - sum(d)^2/n
    st4(ps, make_float4(s.x + n * shift.x, s.y + n * shift.y, s.z + n * shift.z, s.w + n * shift.w));
    st4(pm, make_float4(q.x - s.x * s.x / n, q.y - s.y * s.y / n, q.z - s.z * s.z / n, q.w - s.w * s.w / n));
    if (blockIdx.y == 0 && threadIdx.x == 0) partials[(long long)tiles * 2 * C + tile] = n;
  }
}

// Scalar-channel variant for C % 4 != 0 (tiny layers only).
__global__ void bn_colstats_scalar_kernel(const float* __restrict__ x, long long M, int C, int rows_per_tile, int tiles,
                                          float* __restrict__ partials) {
  const int tile = blockIdx.x;
  const int c = blockIdx.y * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const long long r0 = (long long)tile * rows_per_tile, r1 = min(M, r0 + rows_per_tile);
  const float shift = x[r0 * C + c];
  float s = 0.f, q = 0.f;
  for (long long r = r0; r < r1; ++r) {
    const float d = x[r * C + c] - shift;
    s += d; q = fmaf(d, d, q);
  }
  const float n = (float)(r1 - r0);
  partials[((long long)tile * 2 + 0) * C + c] = s + n * shift;
  partials[((long long)tile * 2 + 1) * C + c] = q - s * s / n;
  if (c == 0) partials[(long long)tiles * 2 * C + tile] = n;
}

struct Moments { double n, mean, m2; };
__device__ __forceinline__ Moments merge(Moments a, Moments b) {
  if (b.n == 0) return a;
  if (a.n == 0) return b;
  Moments r;
  r.n = a.n + b.n;
  const double d = b.mean - a.mean;
  r.mean = a.mean + d * (b.n / r.n);
  r.m2 = a.m2 + b.m2 + d * d * (a.n * b.n / r.n);
  return r;
}

// ---------------------------------------------------------------------------------------------------------
// One-launch cross-replica exchange for sync-BatchNorm (K27; the reference has neither data parallelism nor a
// cross-replica BatchNorm).  Every replica owns an exchange area per BatchNorm layer and direction, mapped by all peers:
//     u64 words[2 parities][world sources][6][C]  |  u64 counter[ceil(C/8)]
// The protocol is flag-in-data ("LL"): each of a channel's three FP64 values travels as two 8-byte words
// {payload: 32 bits of the value, flag: this call's epoch}.  An aligned 8-byte store is indivisible, so a word whose
// flag equals the epoch carries valid data -- no fence, no separate flag store and no acknowledgement: the exchange
// costs ONE one-way NVLink store latency.  (The first version pushed plain data, fenced at system scope and then raised
// flags with st.release: two NVLink round trips on the critical path of every BatchNorm, 10.7 us per exchange and 1.1 ms
// of the 16.1 ms two-GPU step; profiles/06.)
// Warp 0 of a finalize CTA pushes its channels' words into slot [source = this rank] of every replica through the peer
// mappings, then polls ITS OWN area -- local memory, no NVLink polling -- until every source's words carry the epoch,
// and merges the sources in rank order, so all replicas compute bit-identical statistics.  The epoch is a
// device-resident per-(layer, direction, channel group) call counter stored behind the words: no host bookkeeping, no
// counter kernel, replayable from a CUDA graph.  Two parities are enough: a replica can only be one call ahead of the
// slowest one (it needs that replica's words of call e to finish call e, and that replica pushed them after it had
// finished reading call e-1).  This needs a layer's calls to alternate parity per channel group, so one layer must always
// use the same finalize kernel variant: the host side picks the variant for sync calls from the process-wide switch
// alone, never from the tile count.
struct BnSyncArgs {
  unsigned char* area[kMaxPeers];  // replica r's exchange area of this layer and direction, as mapped here
  int rank, world;                 // world <= 1: no exchange
};

__host__ __device__ inline size_t bn_sync_data_bytes(int world, int C) { return (size_t)2 * world * 6 * C * sizeof(unsigned long long); }

__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// Exchanges (v0, v1, v2) of channel c with every replica; take(a, b, c) receives the three values of source 0, 1, ...
// in rank order.  To be called by all 32 lanes of warp 0; `has` marks the lanes that own a channel.
template <typename Take>
__device__ __forceinline__ void bn_sync_exchange(const BnSyncArgs& s, int lane, int C, int c, bool has, double v0,
                                                 double v1, double v2, Take take) {
  const int g = blockIdx.x;
  const size_t data_bytes = bn_sync_data_bytes(s.world, C);
  unsigned char* mine = s.area[s.rank];
  unsigned long long* ctr = reinterpret_cast<unsigned long long*>(mine + data_bytes) + g;
  const unsigned long long call = *reinterpret_cast<volatile unsigned long long*>(ctr) + 1;
  const unsigned long long flag = (call & 0xffffffffull) << 32;
  const size_t parity_off = (size_t)(call & 1) * s.world * 6 * C;   // in words
  if (has) {
    const unsigned long long b0 = (unsigned long long)__double_as_longlong(v0);
    const unsigned long long b1 = (unsigned long long)__double_as_longlong(v1);
    const unsigned long long b2 = (unsigned long long)__double_as_longlong(v2);
    const unsigned long long w[6] = {flag | (b0 & 0xffffffffull), flag | (b0 >> 32), flag | (b1 & 0xffffffffull),
                                     flag | (b1 >> 32),           flag | (b2 & 0xffffffffull), flag | (b2 >> 32)};
    for (int r = 0; r < s.world; ++r) {
      unsigned long long* d = reinterpret_cast<unsigned long long*>(s.area[r]) + parity_off + (size_t)s.rank * 6 * C + c;
#pragma unroll
      for (int k = 0; k < 6; ++k) st_relaxed_sys_u64(d + (size_t)k * C, w[k]);
    }
    const unsigned long long* src = reinterpret_cast<const unsigned long long*>(mine) + parity_off + c;
    for (int r = 0; r < s.world; ++r) {
      unsigned long long got[6];
      unsigned long long spins = 0;
      bool ok;
      do {
        ok = true;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
          got[k] = ld_relaxed_sys_u64(src + ((size_t)r * 6 + k) * C);
          ok = ok && ((got[k] & 0xffffffff00000000ull) == flag);
        }
        if (!ok && ++spins > (1ull << 28)) __trap();  // a lost replica must not hang the GPU forever
      } while (!ok);
      take(__longlong_as_double((long long)((got[0] & 0xffffffffull) | (got[1] << 32))),
           __longlong_as_double((long long)((got[2] & 0xffffffffull) | (got[3] << 32))),
           __longlong_as_double((long long)((got[4] & 0xffffffffull) | (got[5] << 32))));
    }
  }
  __syncwarp();
  if (lane == 0) *ctr = call;
}

// grid: ceil(C/32); block (32 channels, 32 tile lanes).
// peers: optional cross-replica merge (sync-BN, K27): peer_moments[r] points at replica r's exported
// (n, mean, m2)[C] triple array (double), readable over NVLink peer mapping; n_peers == 0 => local only.
__global__ void __launch_bounds__(1024) bn_finalize_kernel(const float* __restrict__ partials, int tiles, int C,
                                                           float eps, float momentum, float* __restrict__ running_mean,
                                                           float* __restrict__ running_var,
                                                           const float* __restrict__ gamma,
                                                           const float* __restrict__ beta, float* __restrict__ mean_out,
                                                           float* __restrict__ invstd_out, float* __restrict__ scale_out,
                                                           float* __restrict__ shift_out,
                                                           const __grid_constant__ BnSyncArgs sync) {
  __shared__ double s_n[32][33], s_s[32][33], s_q[32][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  // Per tile we have (n, sum, M2 about the tile mean).  sum(x^2) of the tile = M2 + sum^2/n, so three plain FP64
  // running sums per thread merge any number of tiles without a serial Chan update (no FP64 divide in the loop);
  // the FP64 accumulation keeps the final M2 = Q - S^2/N accurate even when |mean| >> std.
  double an = 0, as = 0, aq = 0;
  if (c < C) {
    const float* counts = partials + (long long)tiles * 2 * C;
#pragma unroll 4
    for (int t = threadIdx.y; t < tiles; t += 32) {
      const float n = counts[t];
      const float sf = partials[((long long)t * 2 + 0) * C + c];
      const float m2 = partials[((long long)t * 2 + 1) * C + c];
      const float rn = n > 0.f ? 1.f / n : 0.f;
      an += (double)n;
      as += (double)sf;
      aq += (double)m2 + (double)sf * (double)sf * (double)rn;
    }
  }
  s_n[threadIdx.y][threadIdx.x] = an;
  s_s[threadIdx.y][threadIdx.x] = as;
  s_q[threadIdx.y][threadIdx.x] = aq;
  __syncthreads();
  if (threadIdx.y != 0) return;   // warp 0 finishes
  const bool has = c < C;
  for (int i = 1; i < 32; ++i) {
    an += s_n[i][threadIdx.x]; as += s_s[i][threadIdx.x]; aq += s_q[i][threadIdx.x];
  }
  Moments acc{an, an > 0 ? as / an : 0.0, 0.0};
  acc.m2 = an > 0 ? fmax(aq - as * as / an, 0.0) : 0.0;
  if (sync.world > 1) {
    const Moments local = acc;
    acc = Moments{0, 0, 0};
    bn_sync_exchange(sync, threadIdx.x, C, c, has, local.n, local.mean, local.m2,
                     [&](double n, double mean, double m2) { acc = merge(acc, Moments{n, mean, m2}); });
  }
  if (has) {
    const double var = acc.n > 0 ? acc.m2 / acc.n : 0.0;  // biased: used to normalise
    const float invstd = (float)(1.0 / sqrt(var + (double)eps));
    const float mean = (float)acc.mean;
    mean_out[c] = mean;
    invstd_out[c] = invstd;
    const float sc = gamma ? gamma[c] * invstd : invstd;
    scale_out[c] = sc;
    shift_out[c] = (beta ? beta[c] : 0.f) - mean * sc;
    if (running_mean != nullptr) {
      const double unbiased = acc.n > 1 ? acc.m2 / (acc.n - 1) : var;
      running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
      running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
    }
  }
}

// Narrow variants of the two finalize kernels: 8 channels x 128 tile lanes per CTA, i.e. 4x as many CTAs.  The wide
// kernels put a 4 MB partials array (2048 tiles x 256 channels) on 8 SMs and take ~12 us; the work is the same, only
// the mapping changes.  32-byte row segments keep every DRAM sector fully used.  Used for >= 256 tiles
// (DCNN_BN_FINALIZE_NARROW=0 switches them off).
__global__ void __launch_bounds__(1024) bn_finalize_narrow_kernel(const float* __restrict__ partials, int tiles, int C,
                                                                  float eps, float momentum,
                                                                  float* __restrict__ running_mean,
                                                                  float* __restrict__ running_var,
                                                                  const float* __restrict__ gamma,
                                                                  const float* __restrict__ beta,
                                                                  float* __restrict__ mean_out,
                                                                  float* __restrict__ invstd_out,
                                                                  float* __restrict__ scale_out,
                                                                  float* __restrict__ shift_out,
                                                                  const __grid_constant__ BnSyncArgs sync) {
  __shared__ double s_n[128][9], s_s[128][9], s_q[128][9];
  const int c = blockIdx.x * 8 + threadIdx.x;
  double an = 0, as = 0, aq = 0;
  if (c < C) {
    const float* counts = partials + (long long)tiles * 2 * C;
#pragma unroll 4
    for (int t = threadIdx.y; t < tiles; t += 128) {
      const float n = counts[t];
      const float sf = partials[((long long)t * 2 + 0) * C + c];
      const float m2 = partials[((long long)t * 2 + 1) * C + c];
      const float rn = n > 0.f ? 1.f / n : 0.f;
      an += (double)n;
      as += (double)sf;
      aq += (double)m2 + (double)sf * (double)sf * (double)rn;
    }
  }
  s_n[threadIdx.y][threadIdx.x] = an;
  s_s[threadIdx.y][threadIdx.x] = as;
  s_q[threadIdx.y][threadIdx.x] = aq;
  __syncthreads();
  if (threadIdx.y >= 4) return;   // warp 0 (4 rows of 8 threads) finishes; row 0 owns the channels
  const bool has = threadIdx.y == 0 && c < C;
  if (has) {
    for (int i = 1; i < 128; ++i) {
      an += s_n[i][threadIdx.x]; as += s_s[i][threadIdx.x]; aq += s_q[i][threadIdx.x];
    }
  }
  double mean_d = an > 0 ? as / an : 0.0;
  double m2_d = an > 0 ? fmax(aq - as * as / an, 0.0) : 0.0;
  if (sync.world > 1) {
    Moments acc{0, 0, 0};
    bn_sync_exchange(sync, threadIdx.y * 8 + threadIdx.x, C, c, has, an, mean_d, m2_d,
                     [&](double n, double mean, double m2) { acc = merge(acc, Moments{n, mean, m2}); });
    an = acc.n; mean_d = acc.mean; m2_d = acc.m2;
  }
  if (has) {
    const double var = an > 0 ? m2_d / an : 0.0;
    const float invstd = (float)(1.0 / sqrt(var + (double)eps));
    const float mean = (float)mean_d;
    mean_out[c] = mean;
    invstd_out[c] = invstd;
    const float sc = gamma ? gamma[c] * invstd : invstd;
    scale_out[c] = sc;
    shift_out[c] = (beta ? beta[c] : 0.f) - mean * sc;
    if (running_mean != nullptr) {
      const double unbiased = an > 1 ? m2_d / (an - 1) : var;
      running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
      running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
    }
  }
}

__global__ void __launch_bounds__(1024) bn_bwd_finalize_narrow_kernel(const float* __restrict__ partials, int tiles,
                                                                      int C, double count,
                                                                      const float* __restrict__ mean,
                                                                      const float* __restrict__ invstd,
                                                                      const float* __restrict__ gamma,
                                                                      float* __restrict__ dgamma,
                                                                      float* __restrict__ dbeta, float* __restrict__ k1,
                                                                      float* __restrict__ k2, float* __restrict__ k3,
                                                                      const __grid_constant__ BnSyncArgs sync) {
  __shared__ double s_a[128][9], s_b[128][9];
  const int c = blockIdx.x * 8 + threadIdx.x;
  double a = 0, b = 0;
  if (c < C) {
#pragma unroll 8
    for (int t = threadIdx.y; t < tiles; t += 128) {
      a += partials[((long long)t * 2 + 0) * C + c];
      b += partials[((long long)t * 2 + 1) * C + c];
    }
  }
  s_a[threadIdx.y][threadIdx.x] = a;
  s_b[threadIdx.y][threadIdx.x] = b;
  __syncthreads();
  if (threadIdx.y >= 4) return;   // warp 0 finishes; row 0 owns the channels
  const bool has = threadIdx.y == 0 && c < C;
  if (has) {
    for (int i = 1; i < 128; ++i) { a += s_a[i][threadIdx.x]; b += s_b[i][threadIdx.x]; }
    if (dgamma != nullptr) {
      dgamma[c] += (float)b;
      dbeta[c] += (float)a;
    }
  }
  if (sync.world > 1) {
    const double la = a, lb = b, lc = count;
    a = b = count = 0;
    bn_sync_exchange(sync, threadIdx.y * 8 + threadIdx.x, C, c, has, la, lb, lc, [&](double x, double y, double z) { a += x; b += y; count += z; });
  }
  if (has) {
    const float sc = gamma ? gamma[c] * invstd[c] : invstd[c];
    const float ma = (float)(a / count), mb = (float)(b / count);
    k1[c] = sc;
    k2[c] = -invstd[c] * mb * sc;
    k3[c] = (mean[c] * invstd[c] * mb - ma) * sc;
  }
}

// Eval-mode scale/shift from running statistics.
__global__ void bn_eval_coeffs_kernel(const float* __restrict__ rm, const float* __restrict__ rv,
                                      const float* __restrict__ gamma, const float* __restrict__ beta, float eps, int C,
                                      float* __restrict__ scale, float* __restrict__ shift) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float invstd = rsqrtf(rv[c] + eps);
  const float sc = gamma ? gamma[c] * invstd : invstd;
  scale[c] = sc;
  shift[c] = (beta ? beta[c] : 0.f) - rm[c] * sc;
}

// y = x*scale[c] + shift[c] (+ residual) (ReLU).  float4 over channels; C % 4 == 0.
// `mask` (optional, RELU only): one byte per float4, bit j set iff lane j of the float4 passed the ReLU -- 1/16 of
// the tensor instead of the whole output when BatchNorm backward needs the mask of a (BN + residual -> ReLU) unit.
// Element-wise BatchNorm passes: float4 items per thread and loop iteration, and the channel of item i.
constexpr int kEwUnroll = 4;
__device__ __forceinline__ int ew_channel(long long i, int C4) {
  return i <= 0xffffffffll ? (int)((unsigned)i % (unsigned)C4) : (int)(i % C4);
}
static inline int ew_grid_unrolled(long long total4) {
  return ew_grid((total4 + kEwUnroll - 1) / kEwUnroll, 256);
}

template <bool RES, bool RELU>
__global__ void __launch_bounds__(256) bn_apply_kernel(const float* __restrict__ x, const float* __restrict__ scale,
                                                       const float* __restrict__ shift, const float* __restrict__ res,
                                                       float* __restrict__ y, long long total4, int C4,
                                                       uint8_t* __restrict__ mask = nullptr) {
  // kEwUnroll float4 per thread and iteration, all loads issued before the first use: one load in flight per thread
  // left these passes latency-bound at ~5.0 of the 6.55 TB/s copy peak (ncu: 57 % of the warp slots, 32 % issue).
  // The channel index uses 32-bit arithmetic (the 64-bit remainder was a ~50-instruction call per element).
  const long long stride = (long long)gridDim.x * (256 * kEwUnroll);
  for (long long base = blockIdx.x * (long long)(256 * kEwUnroll) + threadIdx.x; base < total4; base += stride) {
    float4 v[kEwUnroll], r[kEwUnroll];
#pragma unroll
    for (int k = 0; k < kEwUnroll; ++k) {
      const long long i = base + k * 256;
      if (i < total4) {
        v[k] = ldcs4(x + i * 4);
        if (RES) r[k] = ldcs4(res + i * 4);
      }
    }
#pragma unroll
    for (int k = 0; k < kEwUnroll; ++k) {
      const long long i = base + k * 256;
      if (i >= total4) break;
      const int c = ew_channel(i, C4) * 4;
      const float4 sc = ldg4(scale + c), sf = ldg4(shift + c);
      float4 o;
      o.x = fmaf(v[k].x, sc.x, sf.x); o.y = fmaf(v[k].y, sc.y, sf.y);
      o.z = fmaf(v[k].z, sc.z, sf.z); o.w = fmaf(v[k].w, sc.w, sf.w);
      if (RES) { o.x += r[k].x; o.y += r[k].y; o.z += r[k].z; o.w += r[k].w; }
      if (RELU) {
        if (mask != nullptr)
          mask[i] = (uint8_t)((o.x > 0.f ? 1 : 0) | (o.y > 0.f ? 2 : 0) | (o.z > 0.f ? 4 : 0) | (o.w > 0.f ? 8 : 0));
        o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f);
      }
      st4(y + i * 4, o);
    }
  }
}

__global__ void bn_apply_scalar_kernel(const float* __restrict__ x, const float* __restrict__ scale,
                                       const float* __restrict__ shift, const float* __restrict__ res,
                                       float* __restrict__ y, long long total, int C, int relu) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    float o = fmaf(x[i], scale[c], shift[c]);
    if (res) o += res[i];
    if (relu) o = fmaxf(o, 0.f);
    y[i] = o;
  }
}

// Backward reduce: per row tile, per channel: (sum g, sum g*xhat) with g = dout * (out > 0 if RELU).
// Same grid/block shape as bn_colstats_kernel; partials [tiles][2][C].
// RELU: 0 = no mask, 1 = mask from the saved activation output (needed when a residual was added before the ReLU),
// 2 = mask recomputed as fmaf(x, scale, shift) > 0 -- bit-identical to what bn_apply_kernel evaluated in the forward
// pass, and one activation-sized read cheaper (`out` then carries scale[C] followed by shift[C]).
// TX = threads along the channel dimension (4 channels each); the other 256 / TX threads stride over the rows.  Narrow
// tensors (C = 64: 16 float4 columns) would leave half of a 32-wide row idle, so TX follows C.
template <int RELU, int TX>
__global__ void __launch_bounds__(256) bn_bwd_reduce_kernel(const float* __restrict__ dout, const float* __restrict__ x,
                                                            const float* __restrict__ out,
                                                            const float* __restrict__ mean,
                                                            const float* __restrict__ invstd, long long M, int C,
                                                            int rows_per_tile, float* __restrict__ partials) {
  constexpr int TY = 256 / TX;
  __shared__ float4 s_a[TY][TX];
  __shared__ float4 s_b[TY][TX];
  // column chunk fastest: CTAs that are resident together read neighbouring 4*TX-channel pieces of the SAME rows, so
  // between them they consume whole rows (whole DRAM pages) instead of a 512-byte piece of every row each
  const int tile = blockIdx.y;
  const int c = (blockIdx.x * TX + threadIdx.x) * 4;
  const long long r0 = (long long)tile * rows_per_tile, r1 = min(M, r0 + rows_per_tile);
  const bool active = c < C;
  float4 a = make_float4(0, 0, 0, 0), b = a;
  if (active) {
    const float4 mu = ldg4(mean + c), is = ldg4(invstd + c);
    float4 sc = make_float4(0, 0, 0, 0), sf = sc;
    if (RELU == 2) { sc = ldg4(out + c); sf = ldg4(out + C + c); }
#pragma unroll 4
    for (long long r = r0 + threadIdx.y; r < r1; r += TY) {
      float4 g = ldg4(dout + r * C + c);
      const float4 v = ldg4(x + r * C + c);
      if (RELU == 1) {
        const float4 o = ldg4(out + r * C + c);
        g.x = o.x > 0.f ? g.x : 0.f; g.y = o.y > 0.f ? g.y : 0.f;
        g.z = o.z > 0.f ? g.z : 0.f; g.w = o.w > 0.f ? g.w : 0.f;
      } else if (RELU == 2) {
        g.x = fmaf(v.x, sc.x, sf.x) > 0.f ? g.x : 0.f; g.y = fmaf(v.y, sc.y, sf.y) > 0.f ? g.y : 0.f;
        g.z = fmaf(v.z, sc.z, sf.z) > 0.f ? g.z : 0.f; g.w = fmaf(v.w, sc.w, sf.w) > 0.f ? g.w : 0.f;
      } else if (RELU == 3) {
        const unsigned mk = reinterpret_cast<const uint8_t*>(out)[(r * C + c) >> 2];
        g.x = (mk & 1u) ? g.x : 0.f; g.y = (mk & 2u) ? g.y : 0.f;
        g.z = (mk & 4u) ? g.z : 0.f; g.w = (mk & 8u) ? g.w : 0.f;
      }
      a.x += g.x; a.y += g.y; a.z += g.z; a.w += g.w;
      b.x = fmaf(g.x, (v.x - mu.x) * is.x, b.x); b.y = fmaf(g.y, (v.y - mu.y) * is.y, b.y);
      b.z = fmaf(g.z, (v.z - mu.z) * is.z, b.z); b.w = fmaf(g.w, (v.w - mu.w) * is.w, b.w);
    }
  }
  s_a[threadIdx.y][threadIdx.x] = a;
  s_b[threadIdx.y][threadIdx.x] = b;
  __syncthreads();
  if (threadIdx.y == 0 && active) {
    for (int i = 1; i < TY; ++i) {
      const float4 p = s_a[i][threadIdx.x], q = s_b[i][threadIdx.x];
      a.x += p.x; a.y += p.y; a.z += p.z; a.w += p.w;
      b.x += q.x; b.y += q.y; b.z += q.z; b.w += q.w;
    }
    st4(partials + ((long long)tile * 2 + 0) * C + c, a);
    st4(partials + ((long long)tile * 2 + 1) * C + c, b);
  }
}

__global__ void bn_bwd_reduce_scalar_kernel(const float* __restrict__ dout, const float* __restrict__ x,
                                            const float* __restrict__ out, const float* __restrict__ mean,
                                            const float* __restrict__ invstd, long long M, int C, int rows_per_tile,
                                            float* __restrict__ partials, int relu) {
  const int tile = blockIdx.x;
  const int c = blockIdx.y * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const long long r0 = (long long)tile * rows_per_tile, r1 = min(M, r0 + rows_per_tile);
  float a = 0.f, b = 0.f;
  for (long long r = r0; r < r1; ++r) {
    float g = dout[r * C + c];
    if (relu == 1 && !(out[r * C + c] > 0.f)) g = 0.f;
    if (relu == 2 && !(fmaf(x[r * C + c], out[c], out[C + c]) > 0.f)) g = 0.f;
    a += g;
    b = fmaf(g, (x[r * C + c] - mean[c]) * invstd[c], b);
  }
  partials[((long long)tile * 2 + 0) * C + c] = a;
  partials[((long long)tile * 2 + 1) * C + c] = b;
}

// Sum backward partials (FP64), accumulate dgamma/dbeta, emit apply coefficients k1,k2,k3.
// Optional cross-replica: export (sum_dy, sum_dy_xhat) and finish in bn_bwd_finalize_sync_kernel.
__global__ void __launch_bounds__(1024) bn_bwd_finalize_kernel(const float* __restrict__ partials, int tiles, int C,
                                                               double count, const float* __restrict__ mean,
                                                               const float* __restrict__ invstd,
                                                               const float* __restrict__ gamma,
                                                               float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                               float* __restrict__ k1, float* __restrict__ k2,
                                                               float* __restrict__ k3,
                                                               const __grid_constant__ BnSyncArgs sync) {
  __shared__ double s_a[32][33], s_b[32][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  double a = 0, b = 0;
  if (c < C) {
#pragma unroll 8
    for (int t = threadIdx.y; t < tiles; t += 32) {
      a += partials[((long long)t * 2 + 0) * C + c];
      b += partials[((long long)t * 2 + 1) * C + c];
    }
  }
  s_a[threadIdx.y][threadIdx.x] = a;
  s_b[threadIdx.y][threadIdx.x] = b;
  __syncthreads();
  if (threadIdx.y != 0) return;   // warp 0 finishes
  const bool has = c < C;
  for (int i = 1; i < 32; ++i) { a += s_a[i][threadIdx.x]; b += s_b[i][threadIdx.x]; }
  if (has && dgamma != nullptr) {  // local contribution; the DP all-reduce sums replicas
    dgamma[c] += (float)b;
    dbeta[c] += (float)a;
  }
  if (sync.world > 1) {
    const double la = a, lb = b, lc = count;
    a = b = count = 0;
    bn_sync_exchange(sync, threadIdx.x, C, c, has, la, lb, lc, [&](double x, double y, double z) { a += x; b += y; count += z; });
  }
  if (has) {
    const float sc = gamma ? gamma[c] * invstd[c] : invstd[c];
    const float ma = (float)(a / count), mb = (float)(b / count);
    // dx = (g - ma - xhat*mb) * sc,  xhat = (x - mean)*invstd
    k1[c] = sc;
    k2[c] = -invstd[c] * mb * sc;
    k3[c] = (mean[c] * invstd[c] * mb - ma) * sc;
  }
}

// dx = g*k1 + x*k2 + k3 with g = dout masked by (out > 0); optionally also writes the masked g (the gradient
// that flows into the residual shortcut).
template <int RELU, bool MASKED>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const float* __restrict__ dout, const float* __restrict__ x,
                                                           const float* __restrict__ out, const float* __restrict__ k1,
                                                           const float* __restrict__ k2, const float* __restrict__ k3,
                                                           float* __restrict__ dx, float* __restrict__ masked,
                                                           long long total4, int C4) {
  // same shape as bn_apply_kernel: kEwUnroll float4 per thread and iteration, loads first
  const long long stride = (long long)gridDim.x * (256 * kEwUnroll);
  for (long long base = blockIdx.x * (long long)(256 * kEwUnroll) + threadIdx.x; base < total4; base += stride) {
    float4 gg[kEwUnroll], vv[kEwUnroll], oo[kEwUnroll];
    unsigned mk[kEwUnroll];
#pragma unroll
    for (int k = 0; k < kEwUnroll; ++k) {
      const long long i = base + k * 256;
      if (i < total4) {
        gg[k] = ldcs4(dout + i * 4);
        vv[k] = ldcs4(x + i * 4);
        if (RELU == 1) oo[k] = ldcs4(out + i * 4);
        if (RELU == 3) mk[k] = reinterpret_cast<const uint8_t*>(out)[i];
      }
    }
#pragma unroll
    for (int k = 0; k < kEwUnroll; ++k) {
      const long long i = base + k * 256;
      if (i >= total4) break;
      const int c = ew_channel(i, C4) * 4;
      float4 g = gg[k];
      const float4 v = vv[k];
      if (RELU == 1) {
        const float4 o = oo[k];
        g.x = o.x > 0.f ? g.x : 0.f; g.y = o.y > 0.f ? g.y : 0.f;
        g.z = o.z > 0.f ? g.z : 0.f; g.w = o.w > 0.f ? g.w : 0.f;
      } else if (RELU == 2) {
        const float4 sc = ldg4(out + c), sf = ldg4(out + C4 * 4 + c);
        g.x = fmaf(v.x, sc.x, sf.x) > 0.f ? g.x : 0.f; g.y = fmaf(v.y, sc.y, sf.y) > 0.f ? g.y : 0.f;
        g.z = fmaf(v.z, sc.z, sf.z) > 0.f ? g.z : 0.f; g.w = fmaf(v.w, sc.w, sf.w) > 0.f ? g.w : 0.f;
      } else if (RELU == 3) {
        g.x = (mk[k] & 1u) ? g.x : 0.f; g.y = (mk[k] & 2u) ? g.y : 0.f;
        g.z = (mk[k] & 4u) ? g.z : 0.f; g.w = (mk[k] & 8u) ? g.w : 0.f;
      }
      const float4 a = ldg4(k1 + c), b = ldg4(k2 + c), d = ldg4(k3 + c);
      float4 r;
      r.x = fmaf(g.x, a.x, fmaf(v.x, b.x, d.x)); r.y = fmaf(g.y, a.y, fmaf(v.y, b.y, d.y));
      r.z = fmaf(g.z, a.z, fmaf(v.z, b.z, d.z)); r.w = fmaf(g.w, a.w, fmaf(v.w, b.w, d.w));
      st4(dx + i * 4, r);
      if (MASKED) st4(masked + i * 4, g);
    }
  }
}

__global__ void bn_bwd_apply_scalar_kernel(const float* __restrict__ dout, const float* __restrict__ x,
                                           const float* __restrict__ out, const float* __restrict__ k1,
                                           const float* __restrict__ k2, const float* __restrict__ k3,
                                           float* __restrict__ dx, float* __restrict__ masked, long long total, int C,
                                           int relu) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    float g = dout[i];
    if (relu == 1 && !(out[i] > 0.f)) g = 0.f;
    if (relu == 2 && !(fmaf(x[i], out[c], out[C + c]) > 0.f)) g = 0.f;
    dx[i] = fmaf(g, k1[c], fmaf(x[i], k2[c], k3[c]));
    if (masked) masked[i] = g;
  }
}

// db[c] += sum over rows of dy[row][c]   (bias gradient; accumulates like cudnnConvolutionBackwardBias beta=1)
__global__ void __launch_bounds__(256) colsum_accum_kernel(const float* __restrict__ dy, long long M, int C,
                                                           int rows_per_block, float* __restrict__ db) {
  __shared__ float s[8][33];
  const int c = blockIdx.y * 32 + threadIdx.x;
  const long long r0 = (long long)blockIdx.x * rows_per_block, r1 = min(M, r0 + rows_per_block);
  float a = 0.f;
  if (c < C)
    for (long long r = r0 + threadIdx.y; r < r1; r += 8) a += __ldg(dy + r * C + c);
  s[threadIdx.y][threadIdx.x] = a;
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    for (int i = 1; i < 8; ++i) a += s[i][threadIdx.x];
    atomicAdd(db + c, a);
  }
}

// ------------------------------------------------------------------------------------------ GroupNorm
// One CTA per (n, group).  The group's HW x Cg slice is read twice (stats, then apply) - the second read hits L2.
__global__ void __launch_bounds__(256) gn_forward_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                         const float* __restrict__ gamma,
                                                         const float* __restrict__ beta, float* __restrict__ mean_out,
                                                         float* __restrict__ invstd_out, int HW, int C, int G,
                                                         float eps) {
  __shared__ float red[2][8];
  __shared__ float s_stat[2];
  const int n = blockIdx.x / G, g = blockIdx.x % G;
  const int Cg = C / G;
  const long long base = (long long)n * HW * C + (long long)g * Cg;
  const int cnt = HW * Cg;
  const float shift = x[base];
  float s = 0.f, q = 0.f;
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    const int p = i / Cg, c = i - p * Cg;
    const float d = x[base + (long long)p * C + c] - shift;
    s += d; q = fmaf(d, d, q);
  }
  s = warp_sum(s); q = warp_sum(q);
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = s; red[1][threadIdx.x >> 5] = q; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, b = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { a += red[0][i]; b += red[1][i]; }
    const float mean_d = a / cnt;
    const float var = fmaxf(b / cnt - mean_d * mean_d, 0.f);
    s_stat[0] = mean_d + shift;
    s_stat[1] = rsqrtf(var + eps);
    mean_out[blockIdx.x] = s_stat[0];
    invstd_out[blockIdx.x] = s_stat[1];
  }
  __syncthreads();
  const float mean = s_stat[0], invstd = s_stat[1];
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    const int p = i / Cg, c = i - p * Cg;
    const long long idx = base + (long long)p * C + c;
    float v = (x[idx] - mean) * invstd;
    if (gamma) v = fmaf(v, gamma[g * Cg + c], beta[g * Cg + c]);
    y[idx] = v;
  }
}

// dgamma/dbeta accumulate (ref groupnorm_ops.cu:175-176 uses +=); dx per group.
__global__ void __launch_bounds__(256) gn_backward_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                          const float* __restrict__ mean,
                                                          const float* __restrict__ invstd,
                                                          const float* __restrict__ gamma, float* __restrict__ dx,
                                                          int HW, int C, int G) {
  __shared__ float red[2][8];
  __shared__ float s_stat[2];
  const int n = blockIdx.x / G, g = blockIdx.x % G;
  const int Cg = C / G;
  const long long base = (long long)n * HW * C + (long long)g * Cg;
  const int cnt = HW * Cg;
  const float mu = mean[blockIdx.x], is = invstd[blockIdx.x];
  float s1 = 0.f, s2 = 0.f;
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    const int p = i / Cg, c = i - p * Cg;
    const long long idx = base + (long long)p * C + c;
    const float gy = dy[idx] * (gamma ? gamma[g * Cg + c] : 1.f);
    s1 += gy;
    s2 = fmaf(gy, (x[idx] - mu) * is, s2);
  }
  s1 = warp_sum(s1); s2 = warp_sum(s2);
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = s1; red[1][threadIdx.x >> 5] = s2; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, b = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { a += red[0][i]; b += red[1][i]; }
    s_stat[0] = a / cnt; s_stat[1] = b / cnt;
  }
  __syncthreads();
  const float m1 = s_stat[0], m2 = s_stat[1];
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    const int p = i / Cg, c = i - p * Cg;
    const long long idx = base + (long long)p * C + c;
    const float gy = dy[idx] * (gamma ? gamma[g * Cg + c] : 1.f);
    const float xh = (x[idx] - mu) * is;
    dx[idx] = (gy - m1 - xh * m2) * is;
  }
}

// dgamma[c] += sum dy*xhat ; dbeta[c] += sum dy   (per channel over N, HW)
__global__ void __launch_bounds__(256) gn_param_grad_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                            const float* __restrict__ mean,
                                                            const float* __restrict__ invstd, float* __restrict__ dgamma,
                                                            float* __restrict__ dbeta, int N, int HW, int C, int G) {
  __shared__ float red[2][8];
  const int c = blockIdx.x;
  const int g = c / (C / G);
  float a = 0.f, b = 0.f;
  const int total = N * HW;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int n = i / HW;
    const long long idx = (long long)i * C + c;
    const float d = dy[idx];
    a += d;
    b = fmaf(d, (x[idx] - mean[n * G + g]) * invstd[n * G + g], b);
  }
  a = warp_sum(a); b = warp_sum(b);
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = a; red[1][threadIdx.x >> 5] = b; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float sa = 0.f, sb = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { sa += red[0][i]; sb += red[1][i]; }
    dbeta[c] += sa;
    dgamma[c] += sb;
  }
}

static int stat_rows_per_tile(long long M, int* tiles) {
  // enough tiles to fill the machine, at least kStatRows rows each
  long long target = (long long)num_sms() * 8;
  long long rows = (M + target - 1) / target;
  if (rows < kStatRows) rows = kStatRows;
  rows = (rows + 7) / 8 * 8;
  *tiles = (int)((M + rows - 1) / rows);
  return (int)rows;
}

}  // namespace dcnn
using namespace dcnn;

DCNN_API int dcnn_bn_stat_tiles(long long M) {
  int tiles;
  stat_rows_per_tile(M, &tiles);
  return tiles;
}

// partials must hold tiles*(2*C+1) floats with tiles = dcnn_bn_stat_tiles(M)
DCNN_API int dcnn_bn_colstats(const float* x, long long M, int C, float* partials, cudaStream_t stream) {
  int tiles;
  const int rows = stat_rows_per_tile(M, &tiles);
  if ((C & 3) == 0) {
    dim3 grid(tiles, ceil_div(C, 128)), block(32, 8);
    bn_colstats_kernel<<<grid, block, 0, stream>>>(x, M, C, rows, tiles, partials);
  } else {
    dim3 grid(tiles, ceil_div(C, 32));
    bn_colstats_scalar_kernel<<<grid, 32, 0, stream>>>(x, M, C, rows, tiles, partials);
  }
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

// read per call (cheap) so that a test can toggle it
static bool finalize_narrow() {
  const char* e = getenv("DCNN_BN_FINALIZE_NARROW");   // default on (-0.34 ms per ResNet-50 step, profiles/05); 0 = wide only
  return e == nullptr || atoi(e) != 0;
}

DCNN_API int dcnn_bn_finalize(const float* partials, int tiles, int C, float eps, float momentum, float* running_mean,
                              float* running_var, const float* gamma, const float* beta, float* mean, float* invstd,
                              float* scale, float* shift, cudaStream_t stream) {
  if (finalize_narrow() && tiles >= 256) {
    dim3 nblock(8, 128);
    bn_finalize_narrow_kernel<<<ceil_div(C, 8), nblock, 0, stream>>>(partials, tiles, C, eps, momentum, running_mean,
                                                                    running_var, gamma, beta, mean, invstd, scale,
                                                                    shift, BnSyncArgs{});
    DCNN_CHECK_LAUNCH();
    DCNN_COUNT_LAUNCH(1);
    return 0;
  }
  dim3 block(32, 32);
  bn_finalize_kernel<<<ceil_div(C, 32), block, 0, stream>>>(partials, tiles, C, eps, momentum, running_mean,
                                                           running_var, gamma, beta, mean, invstd, scale, shift,
                                                           BnSyncArgs{});
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

// ---- one-launch sync-BatchNorm (see bn_sync_exchange): the finalize kernel itself exchanges the per-channel values.
// areas[r]: replica r's exchange area for this layer and direction as mapped in THIS process (host array of `world`
// device pointers); each area holds dcnn_bn_sync_area_bytes(world, C) bytes, zero-initialised once.
DCNN_API long long dcnn_bn_sync_area_bytes(int world, int C) {
  const long long g = (C + 7) / 8;
  const long long b = (long long)bn_sync_data_bytes(world, C) + g * 8;
  return (b + 255) / 256 * 256;
}

static int fill_sync(BnSyncArgs* sa, unsigned char* const* areas, int rank, int world) {
  DCNN_REQUIRE(world >= 2 && world <= kMaxPeers, "sync-BN: %d replicas out of range (2..%d)", world, kMaxPeers);
  DCNN_REQUIRE(rank >= 0 && rank < world, "sync-BN: rank %d out of range", rank);
  for (int r = 0; r < kMaxPeers; ++r) sa->area[r] = r < world ? areas[r] : nullptr;
  sa->rank = rank;
  sa->world = world;
  return 0;
}

// A layer must keep one kernel variant for all of its sync calls (parity alternation, see bn_sync_exchange).
static bool sync_narrow() {
  static const bool on = finalize_narrow();
  return on;
}

DCNN_API int dcnn_bn_finalize_sync1(const float* partials, int tiles, int C, float eps, float momentum,
                                    float* running_mean, float* running_var, const float* gamma, const float* beta,
                                    float* mean, float* invstd, float* scale, float* shift,
                                    unsigned char* const* areas, int rank, int world, cudaStream_t stream) {
  BnSyncArgs sa;
  int rc = fill_sync(&sa, areas, rank, world);
  if (rc) return rc;
  if (sync_narrow()) {
    dim3 nblock(8, 128);
    bn_finalize_narrow_kernel<<<ceil_div(C, 8), nblock, 0, stream>>>(partials, tiles, C, eps, momentum, running_mean,
                                                                    running_var, gamma, beta, mean, invstd, scale,
                                                                    shift, sa);
  } else {
    dim3 block(32, 32);
    bn_finalize_kernel<<<ceil_div(C, 32), block, 0, stream>>>(partials, tiles, C, eps, momentum, running_mean,
                                                             running_var, gamma, beta, mean, invstd, scale, shift,
                                                             sa);
  }
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_bn_bwd_finalize_sync1(const float* partials, int tiles, int C, double count, const float* mean,
                                        const float* invstd, const float* gamma, float* dgamma, float* dbeta,
                                        float* k1, float* k2, float* k3, unsigned char* const* areas, int rank,
                                        int world, cudaStream_t stream) {
  BnSyncArgs sa;
  int rc = fill_sync(&sa, areas, rank, world);
  if (rc) return rc;
  if (sync_narrow()) {
    dim3 nblock(8, 128);
    bn_bwd_finalize_narrow_kernel<<<ceil_div(C, 8), nblock, 0, stream>>>(partials, tiles, C, count, mean, invstd, gamma,
                                                                        dgamma, dbeta, k1, k2, k3, sa);
  } else {
    dim3 block(32, 32);
    bn_bwd_finalize_kernel<<<ceil_div(C, 32), block, 0, stream>>>(partials, tiles, C, count, mean, invstd, gamma,
                                                                 dgamma, dbeta, k1, k2, k3, sa);
  }
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_bn_eval_coeffs(const float* rm, const float* rv, const float* gamma, const float* beta, float eps,
                                 int C, float* scale, float* shift, cudaStream_t stream) {
  bn_eval_coeffs_kernel<<<ceil_div(C, 128), 128, 0, stream>>>(rm, rv, gamma, beta, eps, C, scale, shift);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_bn_apply(const float* x, const float* scale, const float* shift, const float* residual, float* y,
                           long long M, int C, int relu, cudaStream_t stream) {
  const long long total = M * C;
  if ((C & 3) == 0) {
    const long long t4 = total / 4;
    const int grid = ew_grid_unrolled(t4);
    const int C4 = C / 4;
    if (residual && relu) bn_apply_kernel<true, true><<<grid, 256, 0, stream>>>(x, scale, shift, residual, y, t4, C4);
    else if (residual) bn_apply_kernel<true, false><<<grid, 256, 0, stream>>>(x, scale, shift, residual, y, t4, C4);
    else if (relu) bn_apply_kernel<false, true><<<grid, 256, 0, stream>>>(x, scale, shift, residual, y, t4, C4);
    else bn_apply_kernel<false, false><<<grid, 256, 0, stream>>>(x, scale, shift, residual, y, t4, C4);
  } else {
    bn_apply_scalar_kernel<<<ew_grid(total, 256), 256, 0, stream>>>(x, scale, shift, residual, y, total, C, relu);
  }
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

// BN apply + residual/ReLU that also emits the 4-bit-per-float4 ReLU mask (M*C/4 bytes; C % 4 == 0, relu != 0).
// BatchNorm backward then takes the mask through its `out` pointer with relu mode 3.
DCNN_API int dcnn_bn_apply_mask(const float* x, const float* scale, const float* shift, const float* residual,
                                float* y, unsigned char* mask, long long M, int C, cudaStream_t stream) {
  DCNN_REQUIRE((C & 3) == 0, "dcnn_bn_apply_mask: C (%d) must be a multiple of 4", C);
  const long long t4 = M * C / 4;
  const int grid = ew_grid_unrolled(t4);
  if (residual) bn_apply_kernel<true, true><<<grid, 256, 0, stream>>>(x, scale, shift, residual, y, t4, C / 4, mask);
  else bn_apply_kernel<false, true><<<grid, 256, 0, stream>>>(x, scale, shift, residual, y, t4, C / 4, mask);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

// Row tiling of the backward reduce.  It has its own rule (the forward statistics kernel and the allocation-free entry
// point keep stat_rows_per_tile): the tensors it reads are often small (8-70 MB at batch 256), so it wants at least
// four CTAs per SM even when that means 32-row tiles, and it counts the CTAs along the channel dimension.
static int bwd_tx(int C) { return C >= 128 ? 32 : (C >= 64 ? 16 : 8); }
static int bwd_rows_per_tile(long long M, int C, int* tiles) {
  const int tx = bwd_tx(C), ty = 256 / tx;
  const long long col_ctas = (C & 3) == 0 ? ceil_div(C, 4 * tx) : ceil_div(C, 32);
  long long want = ceil_div((long long)num_sms() * 4, col_ctas);
  if (want < 1) want = 1;
  long long rows = ceil_div(M, want);
  if (rows < 32) rows = 32;
  if (rows > 256) rows = 256;
  if (ceil_div(M, rows) > 65535) rows = ceil_div(M, 65535);   // the row tiles are the grid's y dimension
  rows = ceil_div(rows, ty) * ty;
  *tiles = (int)ceil_div(M, rows);
  return (int)rows;
}

// Number of row tiles dcnn_bn_bwd_reduce produces for an [M][C] tensor: its partials buffer holds tiles*2*C floats.
DCNN_API int dcnn_bn_bwd_tiles(long long M, int C) {
  int tiles;
  bwd_rows_per_tile(M, C, &tiles);
  return tiles;
}

template <int TX>
static void launch_bwd_reduce(const float* dout, const float* x, const float* out, const float* mean,
                              const float* invstd, long long M, int C, int relu, int rows, int tiles, float* partials,
                              cudaStream_t stream) {
  dim3 grid(ceil_div(C, 4 * TX), tiles), block(TX, 256 / TX);
  if (relu == 3) bn_bwd_reduce_kernel<3, TX><<<grid, block, 0, stream>>>(dout, x, out, mean, invstd, M, C, rows, partials);
  else if (relu == 2) bn_bwd_reduce_kernel<2, TX><<<grid, block, 0, stream>>>(dout, x, out, mean, invstd, M, C, rows, partials);
  else if (relu) bn_bwd_reduce_kernel<1, TX><<<grid, block, 0, stream>>>(dout, x, out, mean, invstd, M, C, rows, partials);
  else bn_bwd_reduce_kernel<0, TX><<<grid, block, 0, stream>>>(dout, x, out, mean, invstd, M, C, rows, partials);
}

DCNN_API int dcnn_bn_bwd_reduce(const float* dout, const float* x, const float* out, const float* mean,
                                const float* invstd, long long M, int C, int relu, float* partials,
                                cudaStream_t stream) {
  int tiles;
  const int rows = bwd_rows_per_tile(M, C, &tiles);
  if ((C & 3) == 0) {
    const int tx = bwd_tx(C);
    if (tx == 32) launch_bwd_reduce<32>(dout, x, out, mean, invstd, M, C, relu, rows, tiles, partials, stream);
    else if (tx == 16) launch_bwd_reduce<16>(dout, x, out, mean, invstd, M, C, relu, rows, tiles, partials, stream);
    else launch_bwd_reduce<8>(dout, x, out, mean, invstd, M, C, relu, rows, tiles, partials, stream);
  } else {
    dim3 grid(tiles, ceil_div(C, 32));
    bn_bwd_reduce_scalar_kernel<<<grid, 32, 0, stream>>>(dout, x, out, mean, invstd, M, C, rows, partials, relu);
  }
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_bn_bwd_finalize(const float* partials, int tiles, int C, double count, const float* mean,
                                  const float* invstd, const float* gamma, float* dgamma, float* dbeta, float* k1,
                                  float* k2, float* k3, cudaStream_t stream) {
  if (finalize_narrow() && tiles >= 256) {
    dim3 nblock(8, 128);
    bn_bwd_finalize_narrow_kernel<<<ceil_div(C, 8), nblock, 0, stream>>>(partials, tiles, C, count, mean, invstd, gamma,
                                                                        dgamma, dbeta, k1, k2, k3,
                                                                        BnSyncArgs{});
    DCNN_CHECK_LAUNCH();
    DCNN_COUNT_LAUNCH(1);
    return 0;
  }
  dim3 block(32, 32);
  bn_bwd_finalize_kernel<<<ceil_div(C, 32), block, 0, stream>>>(partials, tiles, C, count, mean, invstd, gamma, dgamma,
                                                               dbeta, k1, k2, k3, BnSyncArgs{});
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_bn_bwd_apply(const float* dout, const float* x, const float* out, const float* k1, const float* k2,
                               const float* k3, float* dx, float* masked, long long M, int C, int relu,
                               cudaStream_t stream) {
  const long long total = M * C;
  if ((C & 3) == 0) {
    const long long t4 = total / 4;
    const int grid = ew_grid_unrolled(t4);
    const int C4 = C / 4;
    if (relu == 3 && masked)
      bn_bwd_apply_kernel<3, true><<<grid, 256, 0, stream>>>(dout, x, out, k1, k2, k3, dx, masked, t4, C4);
    else if (relu == 3)
      bn_bwd_apply_kernel<3, false><<<grid, 256, 0, stream>>>(dout, x, out, k1, k2, k3, dx, masked, t4, C4);
    else if (relu == 2 && masked)
      bn_bwd_apply_kernel<2, true><<<grid, 256, 0, stream>>>(dout, x, out, k1, k2, k3, dx, masked, t4, C4);
    else if (relu == 2)
      bn_bwd_apply_kernel<2, false><<<grid, 256, 0, stream>>>(dout, x, out, k1, k2, k3, dx, masked, t4, C4);
    else if (relu && masked)
      bn_bwd_apply_kernel<1, true><<<grid, 256, 0, stream>>>(dout, x, out, k1, k2, k3, dx, masked, t4, C4);
    else if (relu)
      bn_bwd_apply_kernel<1, false><<<grid, 256, 0, stream>>>(dout, x, out, k1, k2, k3, dx, masked, t4, C4);
    else if (masked)
      bn_bwd_apply_kernel<0, true><<<grid, 256, 0, stream>>>(dout, x, out, k1, k2, k3, dx, masked, t4, C4);
    else
      bn_bwd_apply_kernel<0, false><<<grid, 256, 0, stream>>>(dout, x, out, k1, k2, k3, dx, masked, t4, C4);
  } else {
    bn_bwd_apply_scalar_kernel<<<ew_grid(total, 256), 256, 0, stream>>>(dout, x, out, k1, k2, k3, dx, masked, total, C,
                                                                       relu);
  }
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

// deterministic != 0: one CTA per column group walks all rows, so every db[c] receives exactly one atomicAdd per call
// (the sum of two numbers does not depend on the order); otherwise row blocks add their partial sums in arrival order.
DCNN_API int dcnn_colsum_accum(const float* dy, long long M, int C, float* db, int deterministic, cudaStream_t stream) {
  long long target = (long long)num_sms() * 4;
  long long rows = (M + target - 1) / target;
  if (rows < 64) rows = 64;
  if (deterministic) rows = M;
  dim3 grid(ceil_div(M, rows), ceil_div(C, 32)), block(32, 8);
  colsum_accum_kernel<<<grid, block, 0, stream>>>(dy, M, C, (int)rows, db);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_gn_forward(const float* x, float* y, const float* gamma, const float* beta, float* mean,
                             float* invstd, int N, int HW, int C, int G, float eps, cudaStream_t stream) {
  gn_forward_kernel<<<N * G, 256, 0, stream>>>(x, y, gamma, beta, mean, invstd, HW, C, G, eps);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_gn_backward(const float* dy, const float* x, const float* mean, const float* invstd,
                              const float* gamma, float* dgamma, float* dbeta, float* dx, int N, int HW, int C, int G,
                              cudaStream_t stream) {
  if (dgamma != nullptr) {
    gn_param_grad_kernel<<<C, 256, 0, stream>>>(dy, x, mean, invstd, dgamma, dbeta, N, HW, C, G);
    DCNN_CHECK_LAUNCH();
    DCNN_COUNT_LAUNCH(1);
  }
  gn_backward_kernel<<<N * G, 256, 0, stream>>>(dy, x, mean, invstd, gamma, dx, HW, C, G);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Allocation-free BatchNorm batch statistics.
//
// Who needs this: code that is not allowed to call cudaMalloc (or any allocating library routine) at the moment
// it wants BatchNorm statistics -- a host thread inside a latency-critical region that must not take the
// allocator lock, a plugin loaded into somebody else's process that owns no memory pool, or a caller running
// under a fixed memory budget after start-up.  The regular path (dcnn_bn_colstats + dcnn_bn_finalize) needs a
// caller-provided scratch buffer for the per-tile partial sums; here that scratch lives in a file-scope
// `static __device__` array, so the caller passes only inputs and outputs.
//
// Limits and rules:
//   * the scratch holds kNoAllocScratchFloats floats (4 MiB); every GPU that loads this library keeps those
//     4 MiB resident until the process exits.  A call needs tiles*(2*C+1) floats (tiles = dcnn_bn_stat_tiles(M));
//     larger problems are rejected with an error, nothing is launched;
//   * the array name is only a host-side stand-in in host code: the device address is fetched once per device
//     with cudaGetSymbolAddress (return code checked) while that device is current, and that pointer is what the
//     kernels receive.  The first call on each device performs this set-up -- make that first call before any
//     stream capture begins on that device;
//   * all calls share the one array and it keeps the previous call's contents.  Nothing here relies on it being
//     zero: the statistics kernel overwrites every slot (partials and tile counts) that the finalize kernel of the
//     same call reads, so stale values from an earlier, larger call are never observed;
//   * because the scratch is shared, two calls that overlap on the GPU (different streams, or different host
//     threads on the same device) race on it.  Serialise them (same stream, or an event between them).
// The entry point itself calls no cudaMalloc/cudaFree and no allocating library function.
constexpr long long kNoAllocScratchFloats = 1ll << 20;
static __device__ __align__(256) float g_bn_noalloc_scratch[kNoAllocScratchFloats];

namespace {
constexpr int kMaxDevices = 64;
std::atomic<float*> g_noalloc_ptr[kMaxDevices];   // zero-initialised; host threads may race to fill a slot (same value)
}  // namespace

DCNN_API long long dcnn_bn_noalloc_capacity_floats() { return kNoAllocScratchFloats; }

DCNN_API int dcnn_bn_batch_stats_noalloc(const float* x, long long M, int C, float eps, float momentum,
                                         float* running_mean, float* running_var, const float* gamma,
                                         const float* beta, float* mean, float* invstd, float* scale, float* shift,
                                         cudaStream_t stream) {
  DCNN_REQUIRE(M > 0 && C > 0, "dcnn_bn_batch_stats_noalloc: empty input (M=%lld, C=%d)", M, C);
  int tiles;
  const int rows = stat_rows_per_tile(M, &tiles);
  const long long need = (long long)tiles * (2ll * C + 1);
  DCNN_REQUIRE(need <= kNoAllocScratchFloats,
               "dcnn_bn_batch_stats_noalloc: needs %lld scratch floats (M=%lld rows in %d tiles, C=%d) but the built-in "
               "scratch holds %lld; use dcnn_bn_colstats + dcnn_bn_finalize with your own buffer",
               need, M, tiles, C, kNoAllocScratchFloats);
  int dev = 0;
  DCNN_CHECK_CUDA(cudaGetDevice(&dev));
  DCNN_REQUIRE(dev >= 0 && dev < kMaxDevices, "dcnn_bn_batch_stats_noalloc: device index %d out of range", dev);
  float* scratch = g_noalloc_ptr[dev].load(std::memory_order_relaxed);
  if (scratch == nullptr) {  // once per device, with that device current
    void* p = nullptr;
    DCNN_CHECK_CUDA(cudaGetSymbolAddress(&p, g_bn_noalloc_scratch));
    DCNN_REQUIRE(p != nullptr, "dcnn_bn_batch_stats_noalloc: cudaGetSymbolAddress returned a null address");
    scratch = static_cast<float*>(p);
    g_noalloc_ptr[dev].store(scratch, std::memory_order_relaxed);
  }
  if ((C & 3) == 0) {
    dim3 grid(tiles, ceil_div(C, 128)), block(32, 8);
    bn_colstats_kernel<<<grid, block, 0, stream>>>(x, M, C, rows, tiles, scratch);
  } else {
    dim3 grid(tiles, ceil_div(C, 32));
    bn_colstats_scalar_kernel<<<grid, 32, 0, stream>>>(x, M, C, rows, tiles, scratch);
  }
  DCNN_CHECK_LAUNCH();
  dim3 fblock(32, 32);
  bn_finalize_kernel<<<ceil_div(C, 32), fblock, 0, stream>>>(scratch, tiles, C, eps, momentum, running_mean,
                                                            running_var, gamma, beta, mean, invstd, scale, shift,
                                                            BnSyncArgs{});
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(2);
  return 0;
}
