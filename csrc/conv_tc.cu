// Implicit-GEMM convolution / GEMM on the 5th-generation tensor cores (sm_100a).
//
//   * conv_tc{2,3,5}   : D[pixels, oc] = sum_{tap, c} A[pixel@tap, c] * W[oc, tap, c]      (both operands K-major)
//       used for conv fprop, conv dgrad (on dY with re-laid-out weights; strided dgrad as parity classes with an
//       output stride), Dense forward and Dense dgrad.  Replaces cudnnConvolutionForward / BackwardData and
//       cublasSgemm of the reference (src/nn/layers_impl/cuda/cudnn_conv2d_ops.cu:171-216, dense_ops.cu:24-137).
//   * wgrad_tc_kernel  : dW[oc, tap, c] += sum_{pixels} dY[pixel, oc] * X[pixel@tap, c]   (both operands MN-major)
//       replaces cudnnConvolutionBackwardFilter (beta=1 accumulate, cudnn_conv2d_ops.cu:218-232).
//
// Data path: activations are NHWC fp32.  One elected thread issues TMA: the A tile is a 4-D box {32 ch, bw, bh, bn}
// (bw*bh*bn = 128 output pixels) whose start coordinate is shifted by the filter tap; out-of-image coordinates
// (padding) are zero-filled by the TMA unit, so no im2col buffer and no padding copies exist.  Strided convolutions
// use the tensor map's element strides.  Tiles land in shared memory in the 128-byte-swizzled canonical layout and
// are consumed by tcgen05.mma.kind::tf32 (issued by one thread), accumulating FP32 in TMEM.  Four epilogue warps
// read TMEM with tcgen05.ld, add bias, stage the tile in shared memory, then (a) store it coalesced (optionally
// accumulating into the destination, which is how residual-branch gradients are summed) and (b) emit per-channel
// BatchNorm partial statistics (sum, centred M2) for this 128-pixel tile, so BN never re-reads the conv output
// to compute its statistics.
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace dcnn {
using namespace tc;

constexpr int kMaxTaps = 49;
constexpr int kBM = 128;       // output pixels per tile == TMEM lanes
constexpr int kBK = 32;        // fp32 per k-block == 128 bytes == one swizzle row
constexpr int kThreads = 320;   // conv kernels: warp0 TMA, warp1 MMA + TMEM alloc, warps 2-5 and 6-9: two epilogue groups
constexpr int kWgradThreads = 192;  // wgrad: warp0 TMA, warp1 MMA + TMEM alloc, warps 2-5 epilogue

struct ConvTcParams {
  int N, OH, OW;                 // loop-space extents
  int bn, bh, bw;                // tile box (bn*bh*bw == 128)
  int tiles_n, tiles_h, tiles_w;
  int in_mul_h, in_mul_w;        // input coordinate = out coordinate * mul + tap offset
  int Cin, Cout, n_tiles;
  int num_taps;
  int accumulate;
  int relu;                      // epilogue: out = max(out + bias, 0)  (Conv / Dense followed by ReLU without BatchNorm)
  long long out_sn, out_sh, out_sw;
  float* out;
  const float* bias;
  float* partials;               // [m_tiles][2][Cout] then [m_tiles] counts, or nullptr
  int m_tiles;
  int16_t tap_dh[kMaxTaps], tap_dw[kMaxTaps], tap_w[kMaxTaps];
};

// =====================================================================================================
// Epilogue shared by the three kernels below.  TWO groups of four warps (128 threads each; group g uses named barriers
// 1+2g and 2+2g, staging buffers 2g and 2g+1, and takes the 32-column chunks g, g+2, ... of a tile) turn one accumulator
// tile of 128 rows x CHUNKS*32 columns into (a) TMA stores (or TMA reduce-adds) of the output tile and (b) BatchNorm
// partial statistics.  With one group every SM sub-partition ran a single epilogue warp, so nothing hid its LDS / FADD
// latencies (ncu, 1x1 64->256 @32x32: 1600 cycles per chunk for ~450 instructions, tensor pipe 9 %); a tcgen05.ld may
// only touch the 32 TMEM lanes of its warp's quadrant (warp % 4), which warps 6-9 cover again.  The two groups never
// synchronise with each other: each merges the statistics of its own chunks.  Within a group a chunk moves through
// that group's two staging buffers, so that the TMA store of a chunk, the statistics pass over it and the TMEM drain of
// the group's next chunk overlap:
//   per chunk   tcgen05.ld -> (+bias) -> 128B-swizzled staging buffer -> fence.proxy.async -> ONE named barrier
//               -> thread 0: TMA store / reduce-add of that chunk (rows and columns outside the tensor are clipped by
//                  the TMA unit), its own bulk group
//               -> all: shifted single-pass statistics of the chunk (sum(x-k), sum((x-k)^2), k = the tile's first row)
//                  with conflict-free swizzled 64-bit reads: thread (column pair, 16-row group), two-wide FP32 math
//   per tile    the TMEM buffer goes back to the MMA warp right after its last tcgen05.ld; the four row groups are merged
//               through `red` and written as one partial-statistics tile.
// A staging buffer is reused two chunks of the group later; the group's thread 0 lets no bulk group stay unfinished
// before the barrier of a chunk, which frees the buffer of the group's NEXT chunk (cp.async.bulk.wait_group.read 0).
// =====================================================================================================
struct EpiTile {
  int n0, oh0, ow0;     // first output pixel of the tile
  int obn, obh, obw;    // its box (obn * obh * obw == 128 rows, row order n, h, w)
  int col_base;         // first output channel
  int stat_tile;        // index of this tile in the partial-statistics buffer
  int stat_tiles;       // number of statistics tiles (the row counts follow the [stat_tiles][2][Cout] sums)
  bool write_count;     // this tile records its row count (one n-tile per statistics tile does)
  bool full;            // no row of the tile is outside [N, OH, OW]
};

constexpr int kEpiRing = 4;                                  // staging buffers of 16 KB: two per epilogue group
constexpr int kEpiGroups = 2;
constexpr int epi_red_floats(int cols) { return 9 * cols; }  // [2][4 row groups][cols] sums + [cols] shifts

template <int N>
__device__ __forceinline__ void tma_store_wait_read_n() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

// g: epilogue group (0 / 1); e: thread index inside the group (0..127); r: accumulator row = TMEM lane of this thread;
// row_valid: this group's 128 row flags; ring_pos: this group's chunk counter.  release() is called by every thread of
// the group once the group's last tcgen05.ld of the tile has completed.
// POST: the launch has a bias and / or a ReLU to apply (kept out of the instruction stream of the common case -- a
// ResNet convolution has neither -- where the predicated-off loads, adds and maxes were ~140 of the ~300 issue slots
// per chunk of an issue-bound warp).
template <int CHUNKS, bool POST, typename Release>
__device__ __forceinline__ void epi_tile(const ConvTcParams& p, const CUtensorMap* tmO, const EpiTile& t,
                                         uint32_t tmem_d, uint8_t* staging, float* red, uint8_t* row_valid, int g, int r,
                                         int e, uint32_t& ring_pos, Release release) {
  constexpr int RS = CHUNKS * 32;  // row stride of `red`
  const bool stats = p.partials != nullptr;
  if (stats && !t.full) {
    const int w_l = r % t.obw, h_l = (r / t.obw) % t.obh, n_l = r / (t.obw * t.obh);
    row_valid[r] = ((t.n0 + n_l < p.N) && (t.oh0 + h_l < p.OH) && (t.ow0 + w_l < p.OW)) ? 1 : 0;
  }
  const int rg = e >> 5;
  const int bar_a = 1 + 2 * g, bar_b = 2 + 2 * g;
  int cnt = kBM;
  bool any = false, released = false;
#pragma unroll 1
  for (int ch = g; ch < CHUNKS; ch += kEpiGroups) {
    const int c0 = t.col_base + ch * 32;
    if (c0 >= p.Cout) break;
    any = true;
    uint32_t v[32];
    tmem_ld_32x32b_x32(tmem_d + ch * 32, v);
    tmem_ld_wait();
    if (ch + kEpiGroups >= CHUNKS || c0 + 32 * kEpiGroups >= p.Cout) {   // the group's last chunk of this tile
      release();                                                         // accumulator in registers: TMEM back to the MMA warp
      released = true;
    }
    uint8_t* buf = staging + (g * 2 + (ring_pos & 1)) * (kBM * 128);
    uint8_t* dst = buf + r * 128;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 o;
      o.x = __uint_as_float(v[4 * j + 0]); o.y = __uint_as_float(v[4 * j + 1]);
      o.z = __uint_as_float(v[4 * j + 2]); o.w = __uint_as_float(v[4 * j + 3]);
      if constexpr (POST) {
        if (p.bias != nullptr) {
          const int c = c0 + 4 * j;
          if (c < p.Cout) {  // Cout % 4 == 0 on this path
            const float4 b = ldg4(p.bias + c);
            o.x += b.x; o.y += b.y; o.z += b.z; o.w += b.w;
          }
        }
        if (p.relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
      }
      *reinterpret_cast<float4*>(dst + ((j ^ (r & 7)) << 4)) = o;
    }
    // (Writing the chunk with coalesced st.global / red.global from the SMs instead -- on the theory that a TMA store
    // queues behind the mainloop's TMA loads -- was measured and is slower: 15.33 vs 14.23 ms per ResNet-50 step,
    // 1x1 64->256 @32x32 158 vs 97 us; profiles/05 section 7.  The TMA store stays.)
    fence_proxy_async();
    if (e == 0) tma_store_wait_read_n<0>();   // the group's other buffer (its next chunk's) is no longer being read
    asm volatile("bar.sync %0, 128;" ::"r"(bar_a) : "memory");
    if (e == 0) {
      if (p.accumulate) tma_reduce_add_4d(tmO, buf, c0, t.ow0, t.oh0, t.n0);
      else tma_store_4d(tmO, buf, c0, t.ow0, t.oh0, t.n0);
      tma_store_commit();
    }
    ++ring_pos;
    if (stats) {
      if (ch == g && !t.full) {   // the group's first chunk of the tile
        cnt = 0;
        for (int rr = 0; rr < kBM; ++rr) cnt += row_valid[rr];
      }
      // thread (column pair cp, 16-row group hr): 16 LDS.64 and packed two-wide FP32 arithmetic (add / fma .f32x2,
      // SASS FADD2 / FFMA2) -- a third of the issue slots of one column x 32 rows with scalar operations
      const int cp = e & 15, hr = e >> 4;
      const uint32_t cq = cp >> 1, coff = (cp & 1) << 3;
      const uint8_t* rows = buf + hr * 16 * 128;
      const float2 k2 = *reinterpret_cast<const float2*>(buf + (cq << 4) + coff);   // row 0 of the chunk: the shifts
      float2 sum, sq;
      if (t.full) {
        const uint64_t nk = f32x2_pack(-k2.x, -k2.y);
        uint64_t sa = 0, sb = 0, qa = 0, qb = 0;   // the all-zero pattern is {0.f, 0.f}
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
          const uint64_t xa = *reinterpret_cast<const uint64_t*>(rows + i * 128 + ((cq ^ (i & 7)) << 4) + coff);
          const uint64_t xb =
              *reinterpret_cast<const uint64_t*>(rows + (i + 1) * 128 + ((cq ^ ((i + 1) & 7)) << 4) + coff);
          const uint64_t da = f32x2_add(xa, nk), db = f32x2_add(xb, nk);
          sa = f32x2_add(sa, da); qa = f32x2_fma(da, da, qa);
          sb = f32x2_add(sb, db); qb = f32x2_fma(db, db, qb);
        }
        sum = f32x2_unpack(f32x2_add(sa, sb));
        sq = f32x2_unpack(f32x2_add(qa, qb));
      } else {
        sum = make_float2(0.f, 0.f);
        sq = make_float2(0.f, 0.f);
        for (int i = 0; i < 16; ++i) {
          const float2 x = *reinterpret_cast<const float2*>(rows + i * 128 + ((cq ^ (i & 7)) << 4) + coff);
          const bool ok = row_valid[hr * 16 + i] != 0;
          const float d0 = ok ? x.x - k2.x : 0.f, d1 = ok ? x.y - k2.y : 0.f;
          sum.x += d0; sq.x = fmaf(d0, d0, sq.x);
          sum.y += d1; sq.y = fmaf(d1, d1, sq.y);
        }
      }
      // lanes 16-31 hold the second 16 rows of the warp's 32 for the same column pairs
      sum.x += __shfl_xor_sync(0xffffffffu, sum.x, 16); sum.y += __shfl_xor_sync(0xffffffffu, sum.y, 16);
      sq.x += __shfl_xor_sync(0xffffffffu, sq.x, 16); sq.y += __shfl_xor_sync(0xffffffffu, sq.y, 16);
      if ((e & 16) == 0) {
        *reinterpret_cast<float2*>(&red[(0 * 4 + rg) * RS + ch * 32 + 2 * cp]) = sum;
        *reinterpret_cast<float2*>(&red[(1 * 4 + rg) * RS + ch * 32 + 2 * cp]) = sq;
        if (rg == 0) *reinterpret_cast<float2*>(&red[8 * RS + ch * 32 + 2 * cp]) = k2;
      }
    }
  }
  if (!released) release();   // a group without a chunk in this tile (narrow last n-tile) still hands the TMEM buffer back
  if (!stats || !any) return;
  asm volatile("bar.sync %0, 128;" ::"r"(bar_b) : "memory");
  // merge the four row groups of this group's chunks (columns of the chunks g, g+2, ...)
  for (int j = e; j < RS; j += 128) {
    const int c = t.col_base + j;
    if (((j >> 5) % kEpiGroups) != g || c >= p.Cout) continue;   // the other group's chunk, or beyond the last channel
    const float kshift = red[8 * RS + j];
    const float s = red[(0 * 4 + 0) * RS + j] + red[(0 * 4 + 1) * RS + j] + red[(0 * 4 + 2) * RS + j] +
                    red[(0 * 4 + 3) * RS + j];
    const float qq = red[(1 * 4 + 0) * RS + j] + red[(1 * 4 + 1) * RS + j] + red[(1 * 4 + 2) * RS + j] +
                     red[(1 * 4 + 3) * RS + j];
    const float n = (float)cnt;
    p.partials[((long long)t.stat_tile * 2 + 0) * p.Cout + c] = s + n * kshift;
    p.partials[((long long)t.stat_tile * 2 + 1) * p.Cout + c] = cnt > 0 ? fmaxf(qq - s * s / n, 0.f) : 0.f;
  }
  if (e == 0 && g == 0 && t.write_count) p.partials[(long long)t.stat_tiles * 2 * p.Cout + t.stat_tile] = (float)cnt;
}

// =====================================================================================================
// conv_tc2: persistent, warp-specialised implicit GEMM with 128-pixel tiles.
//   * one CTA per SM walks a static tile schedule (n-tile fastest); barriers, TMEM allocation and descriptor
//     prefetch are paid once per CTA instead of once per tile;
//   * the accumulator is double-buffered in TMEM (2 x BN columns): the MMA warp starts tile i+1 while the four
//     epilogue warps drain tile i, so short-K layers (1x1 convs) are bounded by the store stream, not by a serial
//     load -> MMA -> epilogue chain;
//   * the reduction dimension is padded by the TMA unit: a channel count that is a multiple of 4 but not of 32 gets
//     ceil(Cin/32) k-blocks per tap and the out-of-range channels of the A box arrive as zeros (whatever the B box
//     holds there is multiplied by zero);
//   * epilogue: see above.
// =====================================================================================================
template <int BN, int STAGES>
struct ConvTc2Smem {
  static constexpr int kABytes = kBM * 128;
  static constexpr int kBBytes = BN * 128;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kPipeBytes = STAGES * kStageBytes;
  static constexpr int kChunks = BN / 32;
  static constexpr int kStagingBytes = kEpiRing * kBM * 128;
  static constexpr int kRedBytes = epi_red_floats(BN) * 4;
  static constexpr int kBarOff = kPipeBytes + kStagingBytes;
  static constexpr int kRedOff = kBarOff + 256;
  static constexpr int kFlagOff = kRedOff + kRedBytes;
  static constexpr int kTotalBytes = kFlagOff + 256 /*row flags of both epilogue groups*/ + 1024 /*align*/;
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(kThreads, 1) conv_tc2_kernel(const __grid_constant__ CUtensorMap tmA,
                                                               const __grid_constant__ CUtensorMap tmB,
                                                               const __grid_constant__ CUtensorMap tmO,
                                                               const __grid_constant__ ConvTcParams p) {
  using S = ConvTc2Smem<BN, STAGES>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment for the 128B swizzle, computed as an offset so that the compiler keeps the shared address space
  // (a pointer rebuilt from an integer becomes generic: LD.E / ST.E instead of LDS / STS in the epilogue)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* staging = smem + S::kPipeBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarOff);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;   // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;   // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
  float* red = reinterpret_cast<float*>(smem + S::kRedOff);
  uint8_t* row_valid = smem + S::kFlagOff;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_tiles = p.m_tiles * p.n_tiles;
  const int cchunks = (p.Cin + kBK - 1) / kBK;
  const int num_kb = p.num_taps * cchunks;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmO);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < STAGES; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tmem_full_bar[a], 1);
        mbar_init(&tmem_empty_bar[a], 8);  // one arrival per epilogue warp (two groups of four)
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 2 * BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ===================== TMA producer =====================
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int n_tile = tile % p.n_tiles;
        const int m_tile = tile / p.n_tiles;
        const int tw = m_tile % p.tiles_w;
        const int th = (m_tile / p.tiles_w) % p.tiles_h;
        const int tn = m_tile / (p.tiles_w * p.tiles_h);
        const int n0 = tn * p.bn, ih0 = th * p.bh * p.in_mul_h, iw0 = tw * p.bw * p.in_mul_w;
        for (int kb = 0; kb < num_kb; ++kb) {
          const int t = kb / cchunks;
          const int c = kb - t * cchunks;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], S::kStageBytes);
          uint8_t* sA = smem + stage * S::kStageBytes;
          tma_load_4d(sA, &tmA, &full_bar[stage], c * kBK, iw0 + p.tap_dw[t], ih0 + p.tap_dh[t], n0);
          tma_load_2d(sA + S::kABytes, &tmB, &full_bar[stage], p.tap_w[t] * p.Cin + c * kBK, n_tile * BN);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===================== MMA issuer =====================
      constexpr uint32_t idesc = make_idesc_tf32(kBM, BN, 0, 0);
      constexpr uint64_t desc_hi = make_smem_desc_hi(16, 1024);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty_bar[as], aphase ^ 1);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * S::kStageBytes);
          const uint32_t b_addr = a_addr + S::kABytes;
#pragma unroll
          for (int k = 0; k < kBK / 8; ++k)
            mma_tf32_ss(tmem_d, make_smem_desc(a_addr + k * 32, desc_hi), make_smem_desc(b_addr + k * 32, desc_hi),
                        idesc, (kb | k) != 0);
          mma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        mma_commit(&tmem_full_bar[as]);
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int q = warp & 3;          // TMEM lane quadrant
    const int r = q * 32 + lane;     // tile row owned by this thread
    const int g = (warp - 2) >> 2;   // epilogue group: warps 2-5 / 6-9
    const int e = (threadIdx.x - 64) & 127;  // thread index inside the group
    int as = 0;
    uint32_t aphase = 0, ring_pos = 0;
    const bool post = p.bias != nullptr || p.relu != 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int n_tile = tile % p.n_tiles;
      const int m_tile = tile / p.n_tiles;
      const int tw = m_tile % p.tiles_w;
      const int th = (m_tile / p.tiles_w) % p.tiles_h;
      const int tn = m_tile / (p.tiles_w * p.tiles_h);
      EpiTile t;
      t.n0 = tn * p.bn; t.oh0 = th * p.bh; t.ow0 = tw * p.bw;
      t.obn = p.bn; t.obh = p.bh; t.obw = p.bw;
      t.col_base = n_tile * BN;
      t.stat_tile = m_tile; t.stat_tiles = p.m_tiles; t.write_count = (n_tile == 0);
      t.full = (t.n0 + p.bn <= p.N) && (t.oh0 + p.bh <= p.OH) && (t.ow0 + p.bw <= p.OW);

      mbar_wait(&tmem_full_bar[as], aphase);
      tc_fence_after();
      uint64_t* rel_bar = &tmem_empty_bar[as];
      auto release = [&]() {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(rel_bar);
      };
      const uint32_t tmem_d = tmem_base + as * BN + (uint32_t(q * 32) << 16);
      if (post)
        epi_tile<S::kChunks, true>(p, &tmO, t, tmem_d, staging, red, row_valid + g * 128, g, r, e, ring_pos, release);
      else
        epi_tile<S::kChunks, false>(p, &tmO, t, tmem_d, staging, red, row_valid + g * 128, g, r, e, ring_pos, release);
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    if (e == 0) tma_store_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

// =====================================================================================================
// conv_tc3: conv_tc2 with a 256-pixel M tile (two M=128 MMAs per k-step that share one B tile in shared memory).
//
// Why: the ncu capture of conv_tc2 (profiles/03) shows the TF32 mainloop pinned at the L2->SM operand feed
// (~12.5 TB/s chip-wide, 42.6 B/clk/SM), not at the tensor pipe (40 % active) or DRAM (14 %).  A 128x128 tile moves
// (128+128)*128 B per 128x128x32 MAC block = 32 FLOP/B; 256x128 moves (256+128)*128 B for twice the math = 42.7
// FLOP/B, i.e. a 1.33x higher ceiling for every layer regardless of Cout.  The accumulators (2 halves x 2 buffers x BN
// columns) fill TMEM exactly at BN = 128.
//
// The epilogue treats each 128-row half as conv_tc2 treats a tile (own TMA store box, own BN-statistics tile), so
// the partial-statistics buffer has 2 * m_tiles entries.  Default for problems with >= 1 tile per SM (DCNN_CONV_V3=0
// switches it off); measured 408 -> 514 TFLOP/s on the 3x3 128->128 @32x32 layer (profiles/03).
// =====================================================================================================
struct ConvTc3Params {
  ConvTcParams b;     // bn/bh/bw, tiles_* and m_tiles describe the 256-pixel LOAD box
  int obn, obh, obw;  // store / statistics box: one half (128 pixels) of the load box
  int hdn, hdh, hdw;  // coordinate offset of half 1 inside the load box
};

template <int BN, int STAGES>
struct ConvTc3Smem {
  static constexpr int kABytes = 2 * kBM * 128;
  static constexpr int kBBytes = BN * 128;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kPipeBytes = STAGES * kStageBytes;
  static constexpr int kChunks = BN / 32;
  static constexpr int kStagingBytes = kEpiRing * kBM * 128;
  static constexpr int kRedBytes = epi_red_floats(BN) * 4;
  static constexpr int kBarOff = kPipeBytes + kStagingBytes;
  static constexpr int kRedOff = kBarOff + 256;
  static constexpr int kFlagOff = kRedOff + kRedBytes;
  static constexpr int kTotalBytes = kFlagOff + 256 /*row flags of both epilogue groups*/ + 1024 /*align*/;
  static_assert(kTotalBytes <= 232448, "conv_tc3 shared memory exceeds 227 KB");
  static_assert(4 * BN <= 512, "conv_tc3 accumulators exceed TMEM");
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(kThreads, 1) conv_tc3_kernel(const __grid_constant__ CUtensorMap tmA,
                                                               const __grid_constant__ CUtensorMap tmB,
                                                               const __grid_constant__ CUtensorMap tmO,
                                                               const __grid_constant__ ConvTc3Params pp) {
  using S = ConvTc3Smem<BN, STAGES>;
  const ConvTcParams& p = pp.b;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment for the 128B swizzle, computed as an offset so that the compiler keeps the shared address space
  // (a pointer rebuilt from an integer becomes generic: LD.E / ST.E instead of LDS / STS in the epilogue)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* staging = smem + S::kPipeBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarOff);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;   // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;   // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
  float* red = reinterpret_cast<float*>(smem + S::kRedOff);
  uint8_t* row_valid = smem + S::kFlagOff;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_tiles = p.m_tiles * p.n_tiles;
  const int cchunks = (p.Cin + kBK - 1) / kBK;
  const int num_kb = p.num_taps * cchunks;
  constexpr uint32_t kTmemCols = 4 * BN;  // 2 buffers x 2 halves (power of two: 256 or 512)

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmO);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < STAGES; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tmem_full_bar[a], 1);
        mbar_init(&tmem_empty_bar[a], 8);  // one arrival per epilogue warp (two groups of four)
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ===================== TMA producer: one 256-pixel A box + one B box per stage =====================
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int n_tile = tile % p.n_tiles;
        const int m_tile = tile / p.n_tiles;
        const int tw = m_tile % p.tiles_w;
        const int th = (m_tile / p.tiles_w) % p.tiles_h;
        const int tn = m_tile / (p.tiles_w * p.tiles_h);
        const int n0 = tn * p.bn, ih0 = th * p.bh * p.in_mul_h, iw0 = tw * p.bw * p.in_mul_w;
        for (int kb = 0; kb < num_kb; ++kb) {
          const int t = kb / cchunks;
          const int c = kb - t * cchunks;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], S::kStageBytes);
          uint8_t* sA = smem + stage * S::kStageBytes;
          tma_load_4d(sA, &tmA, &full_bar[stage], c * kBK, iw0 + p.tap_dw[t], ih0 + p.tap_dh[t], n0);
          tma_load_2d(sA + S::kABytes, &tmB, &full_bar[stage], p.tap_w[t] * p.Cin + c * kBK, n_tile * BN);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===================== MMA issuer: rows 0-127 -> columns [0,BN), rows 128-255 -> [BN,2BN) =====================
      constexpr uint32_t idesc = make_idesc_tf32(kBM, BN, 0, 0);
      constexpr uint64_t desc_hi = make_smem_desc_hi(16, 1024);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty_bar[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * (2 * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * S::kStageBytes);
          const uint32_t b_addr = a_addr + S::kABytes;
#pragma unroll
          for (int k = 0; k < kBK / 8; ++k) {
            const uint64_t bdesc = make_smem_desc(b_addr + k * 32, desc_hi);
            mma_tf32_ss(tmem_d, make_smem_desc(a_addr + k * 32, desc_hi), bdesc, idesc, (kb | k) != 0);
            mma_tf32_ss(tmem_d + BN, make_smem_desc(a_addr + kBM * 128 + k * 32, desc_hi), bdesc, idesc,
                        (kb | k) != 0);
          }
          mma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        mma_commit(&tmem_full_bar[as]);
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5), one 128-row half at a time =====================
    const int q = warp & 3;
    const int r = q * 32 + lane;     // row inside the half
    const int g = (warp - 2) >> 2;   // epilogue group: warps 2-5 / 6-9
    const int e = (threadIdx.x - 64) & 127;  // thread index inside the group
    int as = 0;
    uint32_t aphase = 0, ring_pos = 0;
    const bool post = p.bias != nullptr || p.relu != 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int n_tile = tile % p.n_tiles;
      const int m_tile = tile / p.n_tiles;
      const int tw = m_tile % p.tiles_w;
      const int th = (m_tile / p.tiles_w) % p.tiles_h;
      const int tn = m_tile / (p.tiles_w * p.tiles_h);

      mbar_wait(&tmem_full_bar[as], aphase);
      tc_fence_after();
#pragma unroll 1
      for (int half = 0; half < 2; ++half) {
        EpiTile t;
        t.n0 = tn * p.bn + half * pp.hdn; t.oh0 = th * p.bh + half * pp.hdh; t.ow0 = tw * p.bw + half * pp.hdw;
        t.obn = pp.obn; t.obh = pp.obh; t.obw = pp.obw;
        t.col_base = n_tile * BN;
        t.stat_tile = m_tile * 2 + half; t.stat_tiles = 2 * p.m_tiles; t.write_count = (n_tile == 0);
        t.full = (t.n0 + pp.obn <= p.N) && (t.oh0 + pp.obh <= p.OH) && (t.ow0 + pp.obw <= p.OW);

        uint64_t* rel_bar = &tmem_empty_bar[as];
        auto release = [&]() {
          if (half == 1) {   // both halves read: hand the accumulator buffer back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(rel_bar);
          }
        };
        const uint32_t tmem_d = tmem_base + as * (2 * BN) + half * BN + (uint32_t(q * 32) << 16);
        if (post)
          epi_tile<S::kChunks, true>(p, &tmO, t, tmem_d, staging, red, row_valid + g * 128, g, r, e, ring_pos, release);
        else
          epi_tile<S::kChunks, false>(p, &tmO, t, tmem_d, staging, red, row_valid + g * 128, g, r, e, ring_pos, release);
      }
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    if (e == 0) tma_store_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// =====================================================================================================
// conv_tc5: CTA-pair version (tcgen05 cta_group::2) for wide layers (Cout >= 256).
//
// Two CTAs of a cluster (one TPC) compute a 256-pixel x 256-channel tile with M = 256 MMAs issued by the leader:
// each CTA loads ITS 128 pixels of A and ITS 128 of the 256 weight rows of B, and owns the accumulator rows of its
// pixels (128 TMEM lanes x 256 columns, double buffered = all 512 columns).  Per CTA that is 128 + 128 operand rows
// per k-step for a 128x256 output slice -- half the TMA traffic per MAC of conv_tc2 and 2/3 of conv_tc3's -- with the
// epilogue of tile i still overlapping the mainloop of tile i+1.
//
// Protocol (after the DeepGEMM / CUTLASS 2-SM pattern): full[s] lives in the leader, count 1 = the leader producer's
// arrive.expect_tx(bytes of both CTAs); both producers' TMA loads count their bytes on it.  The leader's MMA thread frees stage s in BOTH CTAs and signals the accumulator to BOTH epilogues
// with multicast commits.  The epilogue warps of both CTAs (2 x 4) arrive on the leader's tmem_empty[a].
// Validated on hardware (tests/test_gpu_ops.py::test_conv_tc5_cta_pairs); the autotuner (ops/cuda.py) picks it per
// shape where it wins, DCNN_CONV_2CTA=1 forces it wherever it applies.
// =====================================================================================================
template <int STAGES>
struct ConvTc5Smem {
  static constexpr int kBN = 256;                  // tile width (channels) of the pair
  static constexpr int kABytes = kBM * 128;        // this CTA's 128 pixels
  static constexpr int kBBytes = (kBN / 2) * 128;  // this CTA's half of the weight rows
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kPipeBytes = STAGES * kStageBytes;
  static constexpr int kHalfCols = 128;            // the epilogue drains 128 columns at a time
  static constexpr int kChunks = kHalfCols / 32;
  static constexpr int kStagingBytes = kEpiRing * kBM * 128;
  static constexpr int kRedBytes = epi_red_floats(kHalfCols) * 4;
  static constexpr int kBarOff = kPipeBytes + kStagingBytes;
  static constexpr int kRedOff = kBarOff + 256;
  static constexpr int kFlagOff = kRedOff + kRedBytes;
  static constexpr int kTotalBytes = kFlagOff + 256 /*row flags of both epilogue groups*/ + 1024 /*align*/;
  static_assert(kTotalBytes <= 232448, "conv_tc5 shared memory exceeds 227 KB");
};

template <int STAGES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
    conv_tc5_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmO, const __grid_constant__ ConvTcParams p) {
  using S = ConvTc5Smem<STAGES>;
  constexpr int BN = S::kBN, HC = S::kHalfCols;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment for the 128B swizzle, computed as an offset so that the compiler keeps the shared address space
  // (a pointer rebuilt from an integer becomes generic: LD.E / ST.E instead of LDS / STS in the epilogue)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* staging = smem + S::kPipeBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarOff);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;   // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;   // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
  float* red = reinterpret_cast<float*>(smem + S::kRedOff);
  uint8_t* row_valid = smem + S::kFlagOff;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = cluster_ctarank();   // 0 = leader
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int m_pairs = (p.m_tiles + 1) >> 1;
  const int total_tiles = m_pairs * p.n_tiles;   // pair tiles
  const int cchunks = (p.Cin + kBK - 1) / kBK;
  const int num_kb = p.num_taps * cchunks;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmO);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < STAGES; ++s) {
        mbar_init(&full_bar[s], 1);    // (used in the leader) the leader producer's arrive.expect_tx for both CTAs' bytes
        mbar_init(&empty_bar[s], 1);   // one multicast commit
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tmem_full_bar[a], 1);   // one multicast commit
        mbar_init(&tmem_empty_bar[a], 16);  // (used in the leader) 8 epilogue warps of each CTA
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc_pair(tmem_slot, 2 * BN);    // executed by warp 1 of BOTH CTAs
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();   // barriers of both CTAs are initialised before any remote arrive / remote complete_tx
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ===================== TMA producer (both CTAs) =====================
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = pair; tile < total_tiles; tile += num_pairs) {
        const int n_tile = tile % p.n_tiles;
        const int m_tile = (tile / p.n_tiles) * 2 + (int)cta_rank;   // may be == m_tiles (odd count): all out of range
        const int tw = m_tile % p.tiles_w;
        const int th = (m_tile / p.tiles_w) % p.tiles_h;
        const int tn = m_tile / (p.tiles_w * p.tiles_h);
        const int n0 = tn * p.bn, ih0 = th * p.bh * p.in_mul_h, iw0 = tw * p.bw * p.in_mul_w;
        for (int kb = 0; kb < num_kb; ++kb) {
          const int t = kb / cchunks;
          const int c = kb - t * cchunks;
          mbar_wait(&empty_bar[stage], phase ^ 1);   // own slot free (the leader's commit reaches both CTAs)
          const uint32_t leader_full = map_to_cta(smem_u32(&full_bar[stage]), 0);
          uint8_t* sA = smem + stage * S::kStageBytes;
          tma_load_4d_pair(sA, &tmA, leader_full, c * kBK, iw0 + p.tap_dw[t], ih0 + p.tap_dh[t], n0);
          tma_load_2d_pair(sA + S::kABytes, &tmB, leader_full, p.tap_w[t] * p.Cin + c * kBK,
                           n_tile * BN + (int)cta_rank * (BN / 2));
          // Only the leader arrives (expecting the bytes of BOTH CTAs); the peer's loads just count their bytes on the
          // leader's barrier.  A transaction count that runs ahead of the expect_tx is fine -- the phase cannot
          // complete before the leader's arrival -- and the peer cannot run a whole phase ahead because it waits for its
          // own empty barrier, which the leader's MMA commit releases.  (The first version had the peer arrive remotely
          // with release semantics: a MEMBAR + ERRBAR in the peer's producer loop on every k-block, which starved its TMA
          // stream -- tensor pipe 28 % against conv_tc3's 62 % on the same layer, profiles/05.)
          if (cta_rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * S::kStageBytes);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && cta_rank == 0) {
      // ===================== MMA issuer (leader CTA only) =====================
      constexpr uint32_t idesc = make_idesc_tf32(2 * kBM, BN, 0, 0);   // M = 256 over the pair
      constexpr uint64_t desc_hi = make_smem_desc_hi(16, 1024);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int tile = pair; tile < total_tiles; tile += num_pairs) {
        mbar_wait(&tmem_empty_bar[as], aphase ^ 1);   // both epilogues have drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * S::kStageBytes);
          const uint32_t b_addr = a_addr + S::kABytes;
#pragma unroll
          for (int k = 0; k < kBK / 8; ++k)
            mma_tf32_ss_pair(tmem_d, make_smem_desc(a_addr + k * 32, desc_hi), make_smem_desc(b_addr + k * 32, desc_hi),
                             idesc, (kb | k) != 0);
          mma_commit_pair(&empty_bar[stage], 0b11);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        mma_commit_pair(&tmem_full_bar[as], 0b11);
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5 of both CTAs): own 128 pixels, 256 channels in two halves ==========
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const int g = (warp - 2) >> 2;   // epilogue group: warps 2-5 / 6-9
    const int e = (threadIdx.x - 64) & 127;
    int as = 0;
    uint32_t aphase = 0, ring_pos = 0;
    const bool post = p.bias != nullptr || p.relu != 0;
    for (int tile = pair; tile < total_tiles; tile += num_pairs) {
      const int n_tile = tile % p.n_tiles;
      const int m_tile = (tile / p.n_tiles) * 2 + (int)cta_rank;
      const bool live = m_tile < p.m_tiles;   // the odd tile out has no pixels: drain the accumulator, write nothing
      const int tw = m_tile % p.tiles_w;
      const int th = (m_tile / p.tiles_w) % p.tiles_h;
      const int tn = m_tile / (p.tiles_w * p.tiles_h);

      mbar_wait(&tmem_full_bar[as], aphase);
      tc_fence_after();
#pragma unroll 1
      for (int half = 0; half < 2; ++half) {
        EpiTile t;
        t.n0 = tn * p.bn; t.oh0 = th * p.bh; t.ow0 = tw * p.bw;
        t.obn = p.bn; t.obh = p.bh; t.obw = p.bw;
        t.col_base = n_tile * BN + half * HC;
        t.stat_tile = m_tile; t.stat_tiles = p.m_tiles; t.write_count = (n_tile == 0 && half == 0);
        t.full = (t.n0 + p.bn <= p.N) && (t.oh0 + p.bh <= p.OH) && (t.ow0 + p.bw <= p.OW);
        const bool cols_live = live && t.col_base < p.Cout;

        // accumulator fully read by this warp: tell the leader's MMA thread (one arrival per epilogue warp per CTA)
        const uint32_t rel_bar = map_to_cta(smem_u32(&tmem_empty_bar[as]), 0);
        auto release = [&]() {
          if (half == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(rel_bar);
          }
        };
        const uint32_t tmem_d = tmem_base + as * BN + half * HC + (uint32_t(q * 32) << 16);
        if (!cols_live)
          release();
        else if (post)
          epi_tile<S::kChunks, true>(p, &tmO, t, tmem_d, staging, red, row_valid + g * 128, g, r, e, ring_pos, release);
        else
          epi_tile<S::kChunks, false>(p, &tmO, t, tmem_d, staging, red, row_valid + g * 128, g, r, e, ring_pos, release);
      }
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    if (e == 0) tma_store_wait_all();
  }

  // both CTAs are done with each other's barriers, shared memory and the paired tensor memory
  __syncwarp();   // the cluster barrier is .aligned: lanes that skipped the single-lane roles wait for their lane 0 here
  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 2 * BN);
  }
}

// =====================================================================================================
// wgrad
// =====================================================================================================
constexpr int kWPix = 64;  // reduction (pixel) rows per pipeline stage

struct WgradTcParams {
  int N, OH, OW;
  int bn, bh, bw;  // box over dY pixels (product == kWPix)
  int tiles_n, tiles_h, tiles_w;
  int in_mul_h, in_mul_w;
  int Cin, Cout, Ktot;
  int num_taps, oc_tiles, c_tiles, splits;
  int pair;   // 1: two taps per CTA (Cout <= 64, stride 1, same-size output): see dcnn_wgrad_tc
  float* dw;  // [Cout][Ktot]
  float* scratch;  // deterministic mode: [splits][Cout][Ktot] slabs written with plain stores (else nullptr)
  float* db;       // bias gradient [Cout] (+=), or nullptr: column sums of dY by one extra N = 16 MMA against a tile of ones
  // operand descriptors are runtime values so that layout experiments do not need a rebuild
  uint64_t desc_hi_a, desc_hi_b;
  uint32_t idesc, idesc_bias;
  uint32_t k_step_bytes;
  int16_t tap_dh[kMaxTaps], tap_dw[kMaxTaps], tap_w[kMaxTaps];
};

template <int BNW, int STAGES>
struct WgradSmem {
  static constexpr int kABytes = 4 * kWPix * 128;          // 4 boxes of [64 pix][32 oc]
  static constexpr int kBBytes = (BNW / 32) * kWPix * 128;  // BNW/32 boxes of [64 pix][32 c]
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kMainBytes = STAGES * kStageBytes;
  static constexpr int kOnesOff = kMainBytes + 1024;   // [8 pixel rows][128 B] of 1.0f, the B operand of the bias-gradient MMA
  static constexpr int kTotalBytes = kOnesOff + 2048 + 1024 /*align*/;
  static_assert(kTotalBytes <= 232448, "wgrad shared memory exceeds 227 KB");
};

template <int BNW, int STAGES>
__global__ void __launch_bounds__(kWgradThreads) wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmDY,
                                                            const __grid_constant__ CUtensorMap tmX,
                                                            const __grid_constant__ WgradTcParams p) {
  using S = WgradSmem<BNW, STAGES>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment for the 128B swizzle, computed as an offset so that the compiler keeps the shared address space
  // (a pointer rebuilt from an integer becomes generic: LD.E / ST.E instead of LDS / STS in the epilogue)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kMainBytes);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // blockIdx.x -> (c_tile, oc_tile, tap); blockIdx.y -> split of the pixel-tile range
  int b = blockIdx.x;
  const int c_tile = b % p.c_tiles;
  b /= p.c_tiles;
  const int oc_tile = b % p.oc_tiles;
  const int tap = b / p.oc_tiles;
  const int pix_tiles = p.tiles_n * p.tiles_h * p.tiles_w;
  const int per = (pix_tiles + p.splits - 1) / p.splits;
  const int pt_begin = blockIdx.y * per;
  const int pt_end = min(pix_tiles, pt_begin + per);
  const int num_kb = max(0, pt_end - pt_begin);
  const int oc0 = oc_tile * 128, c0 = c_tile * BNW;
  // bias gradient (cudnnConvolutionBackwardBias of the reference, cudnn_conv2d_ops.cu:234-247): db[oc] += sum over the
  // pixels of dY[pixel, oc] = dY^T x ones -- the CTA of the first tap and first channel tile of each oc tile issues one
  // extra N = 16 MMA per k-step against a constant tile of ones (zero-filled out-of-image pixels add nothing)
  const bool bias_cta = p.db != nullptr && tap == 0 && c_tile == 0;
  constexpr uint32_t kTmemCols = 2 * BNW;   // BNW accumulator columns + 16 for the bias sums, next power of two

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmDY);
    tma_prefetch_desc(&tmX);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < STAGES; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      mbar_init(tmem_full_bar, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  } else if (warp >= 2 && bias_cta) {
    float* ones = reinterpret_cast<float*>(smem + S::kOnesOff);
    for (int i = threadIdx.x - 64; i < 512; i += 128) ones[i] = 1.f;
    fence_proxy_async();   // generic-proxy writes become visible to the tensor core's reads of shared memory
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (num_kb > 0) {
    if (warp == 0) {
      if (lane == 0) {
        int stage = 0;
        uint32_t phase = 0;
        const int dh = p.tap_dh[tap], dw_ = p.tap_dw[tap];
        // Only boxes that hold real channels are loaded (a box wholly beyond Cout / Cin would be zero-filled by the TMA
        // unit -- measured: at no cost in time, so this saves nothing but the fill); the accumulator rows / columns of
        // a skipped box hold garbage that the guarded write-back never stores.
        const int na = min(4, (p.Cout - oc0 + 31) >> 5);
        const int nb = min(BNW / 32, (p.Cin - c0 + 31) >> 5);
        // pair mode: rows 0-63 of the tile are dY shifted by -tap(2*tap), rows 64-127 dY shifted by -tap(2*tap+1), and
        // X is loaded unshifted, once for both taps
        const int t1 = 2 * tap + 1;
        const int na_pair = na * (t1 < p.num_taps ? 2 : 1);
        const uint32_t stage_tx = uint32_t((p.pair ? na_pair : na) + nb) * (kWPix * 128);
        for (int kb = 0; kb < num_kb; ++kb) {
          const int pt = pt_begin + kb;
          const int tw = pt % p.tiles_w;
          const int th = (pt / p.tiles_w) % p.tiles_h;
          const int tn = pt / (p.tiles_w * p.tiles_h);
          const int n0 = tn * p.bn, oh0 = th * p.bh, ow0 = tw * p.bw;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], stage_tx);
          uint8_t* sA = smem + stage * S::kStageBytes;
          uint8_t* sB = sA + S::kABytes;
          if (p.pair) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int t = 2 * tap + (i >> 1);
              if (t < p.num_taps && (i & 1) < na)
                tma_load_4d(sA + i * kWPix * 128, &tmDY, &full_bar[stage], (i & 1) * 32, ow0 - p.tap_dw[t],
                            oh0 - p.tap_dh[t], n0);
            }
#pragma unroll
            for (int j = 0; j < BNW / 32; ++j)
              if (j < nb) tma_load_4d(sB + j * kWPix * 128, &tmX, &full_bar[stage], c0 + j * 32, ow0, oh0, n0);
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (i < na) tma_load_4d(sA + i * kWPix * 128, &tmDY, &full_bar[stage], oc0 + i * 32, ow0, oh0, n0);
#pragma unroll
            for (int j = 0; j < BNW / 32; ++j)
              if (j < nb)
                tma_load_4d(sB + j * kWPix * 128, &tmX, &full_bar[stage], c0 + j * 32, ow0 * p.in_mul_w + dw_,
                            oh0 * p.in_mul_h + dh, n0);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        const uint32_t idesc = p.idesc;  // both operands MN-major
        int stage = 0;
        uint32_t phase = 0;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * S::kStageBytes);
          const uint32_t b_addr = a_addr + S::kABytes;
#pragma unroll
          for (int k = 0; k < kWPix / 8; ++k) {
            // advancing K by 8 pixel rows == 8 * 128 B == one swizzle atom
            mma_tf32_ss(tmem_base, make_smem_desc(a_addr + k * p.k_step_bytes, p.desc_hi_a),
                        make_smem_desc(b_addr + k * p.k_step_bytes, p.desc_hi_b), idesc, (kb | k) != 0);
            if (bias_cta)
              mma_tf32_ss(tmem_base + BNW, make_smem_desc(a_addr + k * p.k_step_bytes, p.desc_hi_a),
                          make_smem_desc(smem_u32(smem + S::kOnesOff), p.desc_hi_b), p.idesc_bias, (kb | k) != 0);
          }
          mma_commit(&empty_bar[stage]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        mma_commit(tmem_full_bar);
      }
    } else {
      const int q = warp & 3;
      const int r = q * 32 + lane;  // output row = oc
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
      // pair mode: row r is output channel r % 64 of tap 2*tap + r / 64
      const int etap = p.pair ? 2 * tap + (r >> 6) : tap;
      const int oc = p.pair ? (etap < p.num_taps ? (r & 63) : p.Cout) : oc0 + r;
      if (bias_cta) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + (uint32_t(q * 32) << 16) + BNW, v);   // column 0 of the 16 (identical) sums
        tmem_ld_wait();
        if (oc < p.Cout) atomicAdd(p.db + oc, __uint_as_float(v[0]));
      }
#pragma unroll 1
      for (int cc = 0; cc < BNW; cc += 32) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + (uint32_t(q * 32) << 16) + cc, v);
        tmem_ld_wait();
        if (oc < p.Cout) {
          const long long off = (long long)oc * p.Ktot + (long long)p.tap_w[etap] * p.Cin + c0 + cc;
          if (p.scratch != nullptr) {
            // deterministic mode: this split's slab, summed over the splits in a fixed order by wgrad_reduce_kernel
            float* dst = p.scratch + (long long)blockIdx.y * p.Cout * p.Ktot + off;
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              if (c0 + cc + j < p.Cin)
                *reinterpret_cast<float4*>(dst + j) = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                                                 __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
          } else {
            float* dst = p.dw + off;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              if (c0 + cc + j < p.Cin) {
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j),
                             "f"(__uint_as_float(v[j])), "f"(__uint_as_float(v[j + 1])),
                             "f"(__uint_as_float(v[j + 2])), "f"(__uint_as_float(v[j + 3]))
                             : "memory");
              }
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// dW[i] += sum over the split slabs in index order: the deterministic counterpart of the red.global.add epilogue.
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ scratch, float* __restrict__ dw,
                                                           int splits, long long total4) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    float4 a = *reinterpret_cast<const float4*>(dw + i * 4);
    for (int s = 0; s < splits; ++s) {
      const float4 v = *reinterpret_cast<const float4*>(scratch + ((long long)s * total4 + i) * 4);
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    *reinterpret_cast<float4*>(dw + i * 4) = a;
  }
}

// =====================================================================================================
// host side
// =====================================================================================================
typedef CUresult (*PFN_tmapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                        CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                        CUtensorMapFloatOOBfill);

static PFN_tmapEncodeTiled get_encode() {
  static PFN_tmapEncodeTiled fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_tmapEncodeTiled>(ptr);
  }
  return fn;
}

// NHWC fp32 tensor [N][H][W][C] -> 4-D map, box {32, bw*mul_w, bh*mul_h, bn}, element strides {1,mul_w,mul_h,1}
static int make_tmap_nhwc(CUtensorMap* tm, const float* base, int N, int H, int W, int C, int bn, int bh, int bw,
                          int mul_h, int mul_w, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  PFN_tmapEncodeTiled enc = get_encode();
  DCNN_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t gstr[3] = {(cuuint64_t)C * 4, (cuuint64_t)W * C * 4, (cuuint64_t)H * W * C * 4};
  cuuint32_t box[4] = {32, (cuuint32_t)(bw * mul_w), (cuuint32_t)(bh * mul_h), (cuuint32_t)bn};
  cuuint32_t estr[4] = {1, (cuuint32_t)mul_w, (cuuint32_t)mul_h, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(base), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DCNN_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(NHWC %dx%dx%dx%d box %d,%d,%d mul %d,%d) failed: %d", N, H,
               W, C, bn, bh, bw, mul_h, mul_w, (int)r);
  return 0;
}

// row-major fp32 matrix [rows][cols] -> 2-D map, box {32, box_rows}
static int make_tmap_2d(CUtensorMap* tm, const float* base, long long rows, long long cols, int box_rows) {
  PFN_tmapEncodeTiled enc = get_encode();
  DCNN_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstr[1] = {(cuuint64_t)cols * 4};
  cuuint32_t box[2] = {32, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DCNN_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(2D %lldx%lld box %d) failed: %d", rows, cols, box_rows,
               (int)r);
  return 0;
}

// Pick the power-of-two box (bn,bh,bw) with bn*bh*bw == pixels that covers [N,OH,OW] with the fewest tiles.
static void choose_box(int N, int OH, int OW, int pixels, int mul_h, int mul_w, int* bn, int* bh, int* bw) {
  long long best = -1;
  for (int w = 1; w <= pixels; w <<= 1) {
    if (w * mul_w > 256) break;
    for (int h = 1; w * h <= pixels; h <<= 1) {
      if (h * mul_h > 256) break;
      const int n = pixels / (w * h);
      if (n > 256) continue;
      const long long tiles = (long long)ceil_div(OW, w) * ceil_div(OH, h) * ceil_div(N, n);
      // prefer fewer tiles, then wider rows (longer contiguous runs)
      if (best < 0 || tiles < best || (tiles == best && w > *bw)) {
        best = tiles;
        *bn = n; *bh = h; *bw = w;
      }
    }
  }
}

// Output view [N][OH][OW][Cout] with arbitrary element strides -> 4-D map, box {32, bw, bh, bn}
static int make_tmap_out(CUtensorMap* tm, float* base, int N, int OH, int OW, int Cout, long long sn, long long sh,
                         long long sw, int bn, int bh, int bw) {
  PFN_tmapEncodeTiled enc = get_encode();
  DCNN_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t gdim[4] = {(cuuint64_t)Cout, (cuuint64_t)OW, (cuuint64_t)OH, (cuuint64_t)N};
  cuuint64_t gstr[3] = {(cuuint64_t)sw * 4, (cuuint64_t)sh * 4, (cuuint64_t)sn * 4};
  cuuint32_t box[4] = {32, (cuuint32_t)bw, (cuuint32_t)bh, (cuuint32_t)bn};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DCNN_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(out %dx%dx%dx%d strides %lld,%lld,%lld) failed: %d", N, OH,
               OW, Cout, sn, sh, sw, (int)r);
  return 0;
}

template <int BN, int STAGES>
static int launch_conv2(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmO, const ConvTcParams& p,
                        cudaStream_t stream) {
  using S = ConvTc2Smem<BN, STAGES>;
  DCNN_SET_MAX_SMEM_ONCE(S::kTotalBytes, conv_tc2_kernel<BN, STAGES>);
  const long long tiles = (long long)p.m_tiles * p.n_tiles;
  const int grid = (int)(tiles < num_sms() ? tiles : num_sms());
  conv_tc2_kernel<BN, STAGES><<<grid, kThreads, S::kTotalBytes, stream>>>(tmA, tmB, tmO, p);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

template <int BN, int STAGES>
static int launch_conv3(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmO,
                        const ConvTc3Params& p, cudaStream_t stream) {
  using S = ConvTc3Smem<BN, STAGES>;
  DCNN_SET_MAX_SMEM_ONCE(S::kTotalBytes, conv_tc3_kernel<BN, STAGES>);
  const long long tiles = (long long)p.b.m_tiles * p.b.n_tiles;
  const int grid = (int)(tiles < num_sms() ? tiles : num_sms());
  conv_tc3_kernel<BN, STAGES><<<grid, kThreads, S::kTotalBytes, stream>>>(tmA, tmB, tmO, p);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

template <int STAGES>
static int launch_conv5(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmO, const ConvTcParams& p,
                        cudaStream_t stream) {
  using S = ConvTc5Smem<STAGES>;
  DCNN_SET_MAX_SMEM_ONCE(S::kTotalBytes, conv_tc5_kernel<STAGES>);
  const long long pair_tiles = (long long)((p.m_tiles + 1) / 2) * p.n_tiles;
  const int max_pairs = num_sms() / 2;
  const int pairs = (int)(pair_tiles < max_pairs ? pair_tiles : max_pairs);
  conv_tc5_kernel<STAGES><<<2 * pairs, kThreads, S::kTotalBytes, stream>>>(tmA, tmB, tmO, p);  // clusters of 2
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Which kernel runs a problem.  All three are persistent (one CTA, or CTA pair, per SM, static round-robin) and bound by
// the L2 -> SM operand feed once K is deep, so a launch costs about  waves x k-blocks x (time of one k-block), with the
// k-block times measured on B200 (profiles/05 section 3 and section 5):
//     conv_tc2  128 px x 128 ch per CTA   256 operand rows   0.39 us     (N = 64: 192 rows)
//     conv_tc3  256 px x 128 ch per CTA   384 rows           0.53 us     (N = 64: 320 rows)
//     conv_tc5  128 px x 256 ch per CTA   256 rows           0.45 us     (a CTA pair shares one 256 x 256 tile)
// conv_tc5 does twice conv_tc2's work per k-block for 1.15x its time, but its cross-CTA handshakes and its two-pass
// epilogue only pay off when the mainloop is long: it needs Cout >= 256 and at least 16 k-blocks.
// Per-thread override, used by the autotuner (ops/cuda.py) and the tests: 0 = this rule, 2 / 3 / 5 = that kernel wherever
// it is structurally possible.  DCNN_CONV_2CTA=1 forces conv_tc5 wherever it applies, DCNN_CONV_2CTA=0 and DCNN_CONV_V3=0
// take the respective kernel out of the rule.
static thread_local int g_conv_variant = 0;

static bool conv_v3_enabled() {
  static const bool on = !(getenv("DCNN_CONV_V3") != nullptr && atoi(getenv("DCNN_CONV_V3")) == 0);
  return on;
}
static int conv_2cta_mode() {   // -1: not set, 0: off, 1: forced
  static const int m = getenv("DCNN_CONV_2CTA") == nullptr ? -1 : (atoi(getenv("DCNN_CONV_2CTA")) != 0 ? 1 : 0);
  return m;
}

struct ConvChoice {
  int variant;           // 2, 3 or 5
  int bn3, bh3, bw3;     // conv_tc3's 256-pixel load box (valid when variant == 3)
};

static ConvChoice choose_conv_variant(int N, int OH, int OW, int mul_h, int mul_w, int Cout, int num_kb) {
  ConvChoice c;
  c.variant = 2;
  c.bn3 = c.bh3 = c.bw3 = 1;
  choose_box(N, OH, OW, 2 * kBM, mul_h, mul_w, &c.bn3, &c.bh3, &c.bw3);
  const bool can5 = Cout >= 256;
  if (g_conv_variant == 2) return c;
  if (g_conv_variant == 3) { c.variant = 3; return c; }
  if (g_conv_variant == 5) { c.variant = can5 ? 5 : 2; return c; }
  if (conv_2cta_mode() == 1 && can5) { c.variant = 5; return c; }
  int b2n = 1, b2h = 1, b2w = 1;
  choose_box(N, OH, OW, kBM, mul_h, mul_w, &b2n, &b2h, &b2w);
  const long long m2 = (long long)ceil_div(OW, b2w) * ceil_div(OH, b2h) * ceil_div(N, b2n);
  const long long m3 = (long long)ceil_div(OW, c.bw3) * ceil_div(OH, c.bh3) * ceil_div(N, c.bn3);
  const long long sms = num_sms();
  const bool narrow = Cout <= 64;
  const long long n_tiles = narrow ? 1 : ceil_div(Cout, 128);
  // k-block times in units of 0.01 us
  const long long t2 = narrow ? 29 : 39, t3 = narrow ? 44 : 53, t5 = 45;
  long long best = ((m2 * n_tiles + sms - 1) / sms) * t2;
  if (conv_v3_enabled()) {
    const long long cost3 = ((m3 * n_tiles + sms - 1) / sms) * t3;
    if (cost3 < best) { best = cost3; c.variant = 3; }
  }
  if (can5 && num_kb >= 16 && conv_2cta_mode() != 0) {
    const long long pairs = sms / 2;
    const long long cost5 = ((((m2 + 1) / 2) * ceil_div(Cout, 256) + pairs - 1) / pairs) * t5;
    if (cost5 < best) { best = cost5; c.variant = 5; }
  }
  return c;
}

template <int BNW, int STAGES>
static int launch_wgrad(const CUtensorMap& tmDY, const CUtensorMap& tmX, const WgradTcParams& p, cudaStream_t stream) {
  using S = WgradSmem<BNW, STAGES>;
  DCNN_SET_MAX_SMEM_ONCE(S::kTotalBytes, wgrad_tc_kernel<BNW, STAGES>);
  const int tap_groups = p.pair ? (p.num_taps + 1) / 2 : p.num_taps;
  dim3 grid((unsigned)(tap_groups * p.oc_tiles * p.c_tiles), (unsigned)p.splits);
  wgrad_tc_kernel<BNW, STAGES><<<grid, kWgradThreads, S::kTotalBytes, stream>>>(tmDY, tmX, p);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

}  // namespace dcnn

using namespace dcnn;

// Number of 128-pixel M tiles the tensor-core conv uses for a loop space [N,OH,OW]; the BN partial-statistics
// buffer must hold m_tiles*(2*Cout+1) floats.
DCNN_API int dcnn_conv_tc_m_tiles(int N, int OH, int OW, int mul_h, int mul_w) {
  int bn = 1, bh = 1, bw = 1;
  choose_box(N, OH, OW, kBM, mul_h, mul_w, &bn, &bh, &bw);
  return ceil_div(OW, bw) * ceil_div(OH, bh) * ceil_div(N, bn);
}

// Autotuner hook: force a convolution kernel variant for the calls this thread makes next (0 restores the rules).
DCNN_API int dcnn_conv_tc_set_variant(int variant) {
  DCNN_REQUIRE(variant == 0 || variant == 2 || variant == 3 || variant == 5, "unknown conv variant %d", variant);
  g_conv_variant = variant;
  return 0;
}

// Number of BatchNorm partial-statistics tiles dcnn_conv_tc will emit for this problem (the partials buffer holds
// tiles*(2*Cout+1) floats); num_kb = taps * ceil(Cin / 32).  Equals dcnn_conv_tc_m_tiles unless the 256-pixel kernel is
// selected, which emits two 128-row statistics tiles per M tile.
DCNN_API int dcnn_conv_tc_stat_tiles(int N, int OH, int OW, int mul_h, int mul_w, int Cout, int num_kb) {
  const ConvChoice c = choose_conv_variant(N, OH, OW, mul_h, mul_w, Cout, num_kb);
  if (c.variant == 3) return 2 * ceil_div(OW, c.bw3) * ceil_div(OH, c.bh3) * ceil_div(N, c.bn3);
  return dcnn_conv_tc_m_tiles(N, OH, OW, mul_h, mul_w);
}

// Generic tap-list implicit GEMM (see file header).  x: NHWC [N][H][W][Cin] (Cin % 4 == 0: the TMA unit pads each tap's
// channel run to a multiple of 32 with zeros), Cout % 4 == 0, w: [Cout][Ktot] with K index = tap_w*Cin + c.  Output pixel (n,oh,ow), channel oc is written to
// out[n*out_sn + oh*out_sh + ow*out_sw + oc].
DCNN_API int dcnn_conv_tc(const float* x, int N, int H, int W, int Cin, const float* w, int Cout, int Ktot, float* out,
                          long long out_sn, long long out_sh, long long out_sw, int OH, int OW, int in_mul_h,
                          int in_mul_w, int num_taps, const int* tap_dh, const int* tap_dw, const int* tap_w,
                          const float* bias, int accumulate, int relu, float* partials, cudaStream_t stream) {
  DCNN_REQUIRE(!(relu && (accumulate || partials != nullptr)),
               "dcnn_conv_tc: the ReLU epilogue does not combine with accumulation or BatchNorm statistics");
  DCNN_REQUIRE(Cin % 4 == 0 && Ktot % 4 == 0, "dcnn_conv_tc: Cin (%d) and Ktot (%d) must be multiples of 4", Cin, Ktot);
  DCNN_REQUIRE(Cout % 4 == 0, "dcnn_conv_tc: Cout (%d) must be a multiple of 4 (TMA store rows are 16-byte aligned)", Cout);
  DCNN_REQUIRE(num_taps >= 1 && num_taps <= kMaxTaps, "dcnn_conv_tc: bad tap count %d", num_taps);
  DCNN_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(w) & 15) == 0,
               "dcnn_conv_tc: operands must be 16-byte aligned");
  ConvTcParams p;
  memset(&p, 0, sizeof(p));
  p.N = N; p.OH = OH; p.OW = OW;
  choose_box(N, OH, OW, kBM, in_mul_h, in_mul_w, &p.bn, &p.bh, &p.bw);
  p.tiles_n = ceil_div(N, p.bn); p.tiles_h = ceil_div(OH, p.bh); p.tiles_w = ceil_div(OW, p.bw);
  p.m_tiles = p.tiles_n * p.tiles_h * p.tiles_w;
  p.in_mul_h = in_mul_h; p.in_mul_w = in_mul_w;
  p.Cin = Cin; p.Cout = Cout;
  p.num_taps = num_taps; p.accumulate = accumulate; p.relu = relu;
  p.out_sn = out_sn; p.out_sh = out_sh; p.out_sw = out_sw;
  p.out = out; p.bias = bias; p.partials = partials;
  for (int t = 0; t < num_taps; ++t) {
    p.tap_dh[t] = (int16_t)tap_dh[t]; p.tap_dw[t] = (int16_t)tap_dw[t]; p.tap_w[t] = (int16_t)tap_w[t];
  }
  DCNN_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0 && (out_sn & 3) == 0 && (out_sh & 3) == 0 &&
                   (out_sw & 3) == 0,
               "dcnn_conv_tc: output must be 16-byte aligned with strides multiple of 4");
  CUtensorMap tmA, tmB;
  int rc = make_tmap_nhwc(&tmA, x, N, H, W, Cin, p.bn, p.bh, p.bw, in_mul_h, in_mul_w);
  if (rc) return rc;
  const ConvChoice choice = choose_conv_variant(N, OH, OW, in_mul_h, in_mul_w, Cout, num_taps * ceil_div(Cin, kBK));
  if (choice.variant == 5) {
    CUtensorMap tmO5;
    rc = make_tmap_out(&tmO5, out, N, OH, OW, Cout, out_sn, out_sh, out_sw, p.bn, p.bh, p.bw);
    if (rc) return rc;
    p.n_tiles = ceil_div(Cout, 256);
    rc = make_tmap_2d(&tmB, w, Cout, Ktot, 128);   // each CTA of the pair loads 128 of the tile's 256 weight rows
    if (rc) return rc;
    return launch_conv5<4>(tmA, tmB, tmO5, p, stream);
  }
  if (choice.variant == 3) {
    const int bn3 = choice.bn3, bh3 = choice.bh3, bw3 = choice.bw3;
    ConvTc3Params q;
    q.b = p;
    q.b.bn = bn3; q.b.bh = bh3; q.b.bw = bw3;
    q.b.tiles_n = ceil_div(N, bn3); q.b.tiles_h = ceil_div(OH, bh3); q.b.tiles_w = ceil_div(OW, bw3);
    q.b.m_tiles = q.b.tiles_n * q.b.tiles_h * q.b.tiles_w;
    // the second 128 rows of the box (row order n, h, w) start half-way along its slowest dimension of extent >= 2
    q.obn = bn3; q.obh = bh3; q.obw = bw3;
    q.hdn = q.hdh = q.hdw = 0;
    if (bn3 >= 2) { q.obn = bn3 / 2; q.hdn = bn3 / 2; }
    else if (bh3 >= 2) { q.obh = bh3 / 2; q.hdh = bh3 / 2; }
    else { q.obw = bw3 / 2; q.hdw = bw3 / 2; }
    CUtensorMap tmA3, tmO3;
    rc = make_tmap_nhwc(&tmA3, x, N, H, W, Cin, bn3, bh3, bw3, in_mul_h, in_mul_w);
    if (rc) return rc;
    rc = make_tmap_out(&tmO3, out, N, OH, OW, Cout, out_sn, out_sh, out_sw, q.obn, q.obh, q.obw);
    if (rc) return rc;
    if (Cout <= 64) {
      q.b.n_tiles = 1;
      rc = make_tmap_2d(&tmB, w, Cout, Ktot, 64);
      if (rc) return rc;
      return launch_conv3<64, 3>(tmA3, tmB, tmO3, q, stream);
    }
    q.b.n_tiles = ceil_div(Cout, 128);
    rc = make_tmap_2d(&tmB, w, Cout, Ktot, 128);
    if (rc) return rc;
    return launch_conv3<128, 3>(tmA3, tmB, tmO3, q, stream);
  }
  CUtensorMap tmO;
  rc = make_tmap_out(&tmO, out, N, OH, OW, Cout, out_sn, out_sh, out_sw, p.bn, p.bh, p.bw);
  if (rc) return rc;
  if (Cout <= 64) {
    p.n_tiles = 1;
    rc = make_tmap_2d(&tmB, w, Cout, Ktot, 64);
    if (rc) return rc;
    return launch_conv2<64, 6>(tmA, tmB, tmO, p, stream);
  }
  p.n_tiles = ceil_div(Cout, 128);
  rc = make_tmap_2d(&tmB, w, Cout, Ktot, 128);
  if (rc) return rc;
  return launch_conv2<128, 4>(tmA, tmB, tmO, p, stream);
}

// dW[oc][tap_w*Cin + c] += sum_{n,oh,ow} dY[n,oh,ow,oc] * X[n, oh*mul_h + tap_dh, ow*mul_w + tap_dw, c]
// dy: NHWC [N][OH][OW][Cout]; x: NHWC [N][H][W][Cin] (Cin % 4 == 0, Cout % 4 == 0; channel counts that are not
// multiples of 32 are padded with zeros by the TMA unit, the write-back is guarded).
static void wgrad_geometry(int N, int OH, int OW, int Cout, int Cin, int in_mul_h, int in_mul_w, int num_taps,
                           WgradTcParams* p, int* bnw_out) {
  choose_box(N, OH, OW, kWPix, in_mul_h, in_mul_w, &p->bn, &p->bh, &p->bw);
  p->tiles_n = ceil_div(N, p->bn); p->tiles_h = ceil_div(OH, p->bh); p->tiles_w = ceil_div(OW, p->bw);
  p->oc_tiles = ceil_div(Cout, 128);
  // (128 x 256 tiles were tried -- 42.7 instead of 32 FLOP per operand byte, two 96 KB stages -- and measured no faster
  // than 128 x 128 with three stages on the ResNet-50 step: profiles/05.)
  const int bnw = Cin >= 128 ? 128 : 64;
  p->c_tiles = ceil_div(Cin, bnw);
  const int pix_tiles = p->tiles_n * p->tiles_h * p->tiles_w;
  const int base = num_taps * p->oc_tiles * p->c_tiles;
  // Split-K factor.  One CTA fits per SM (its pipeline stages fill shared memory), so a launch runs in
  // ceil(CTAs / SMs) waves and each wave costs the k-iterations of one CTA plus its prologue and red.add epilogue
  // (about four k-iterations).  Pick the factor that minimises waves x (k-iterations + 4); fewer splits on a tie (less
  // atomic traffic).  (Round 1 aimed for two CTAs per SM regardless: 5 splits x 64 CTAs = 2.2 waves for the 1x1
  // 2048->512 layer at 4x4 where 2 splits run in one.)
  const long long sms = num_sms();
  int best = 1;
  long long best_cost = -1;
  const int max_splits = pix_tiles < 512 ? pix_tiles : 512;
  for (int sp = 1; sp <= max_splits; ++sp) {
    const int per = ceil_div(pix_tiles, sp);
    const int eff = ceil_div(pix_tiles, per);   // every split non-empty
    if (eff != sp) continue;
    const long long waves = ((long long)base * eff + sms - 1) / sms;
    const long long cost = waves * (per + 4);
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = eff; }
  }
  p->splits = best;
  *bnw_out = bnw;
}

// Number of split-K slabs dcnn_wgrad_tc uses for this problem: the deterministic mode's scratch buffer holds
// splits * Cout * Ktot floats.
DCNN_API int dcnn_wgrad_tc_splits(int N, int OH, int OW, int Cout, int Cin, int in_mul_h, int in_mul_w, int num_taps) {
  WgradTcParams p;
  memset(&p, 0, sizeof(p));
  int bnw;
  wgrad_geometry(N, OH, OW, Cout, Cin, in_mul_h, in_mul_w, num_taps, &p, &bnw);
  return p.splits;
}

// dW[oc][tap_w*Cin + c] += sum_{n,oh,ow} dY[n,oh,ow,oc] * X[n, oh*mul_h + tap_dh, ow*mul_w + tap_dw, c]
// dy: NHWC [N][OH][OW][Cout]; x: NHWC [N][H][W][Cin] (Cin % 4 == 0, Cout % 4 == 0; channel counts that are not
// multiples of 32 are padded with zeros by the TMA unit, the write-back is guarded).
// db != NULL: the bias gradient db[oc] += sum over pixels of dY[pixel, oc] is computed by the same launch (one extra
// N = 16 MMA per k-step in one CTA per oc tile), replacing the separate column-sum pass.
// scratch == NULL: split-K partial sums are added with red.global.add (fast; the order of the additions, and so the last
// bits of dW, vary from run to run).  scratch != NULL (dcnn_wgrad_tc_splits * Cout * Ktot floats): every split writes
// its own slab and a second kernel sums the slabs in index order -- bit-reproducible gradients.
DCNN_API int dcnn_wgrad_tc(const float* dy, int N, int OH, int OW, int Cout, const float* x, int H, int W, int Cin,
                           float* dw, int Ktot, int in_mul_h, int in_mul_w, int num_taps, const int* tap_dh,
                           const int* tap_dw, const int* tap_w, float* scratch, float* db, cudaStream_t stream) {
  DCNN_REQUIRE(Cin % 4 == 0, "dcnn_wgrad_tc: Cin (%d) must be a multiple of 4", Cin);
  DCNN_REQUIRE(!(db != nullptr && scratch != nullptr),
               "dcnn_wgrad_tc: the fused bias gradient adds split sums in arrival order; use dcnn_colsum_accum in the "
               "deterministic mode");
  DCNN_REQUIRE(Cout % 4 == 0, "dcnn_wgrad_tc: Cout (%d) must be a multiple of 4", Cout);
  DCNN_REQUIRE(num_taps >= 1 && num_taps <= kMaxTaps, "dcnn_wgrad_tc: bad tap count %d", num_taps);
  WgradTcParams p;
  memset(&p, 0, sizeof(p));
  p.N = N; p.OH = OH; p.OW = OW;
  int bnw;
  // Two taps per CTA when the 128-row tile would be half empty (Cout <= 64): with a same-size stride-1 output,
  //   dW[oc][tap][c] = sum over x positions of dY[pos - tap][oc] * X[pos][c]   (dY zero outside the image),
  // so rows 0-63 / 64-127 of the A tile hold dY shifted by two different taps and one unshifted X tile serves both: half
  // the CTAs and MMAs, and 48 KB instead of 64 KB of operand traffic per k-block and tap pair (3x3 64->64 at 32x32,
  // batch 256: 166 -> 98 us, the ResNet-50 step 13.58 -> 13.38 ms; profiles/05 section 8).  Not with the fused bias
  // gradient (its column sums need unshifted dY) nor in the deterministic mode (dcnn_wgrad_tc_splits sizes the scratch
  // from the unpaired geometry).
  p.pair = (Cout <= 64 && num_taps >= 2 && in_mul_h == 1 && in_mul_w == 1 && OH == H && OW == W && db == nullptr &&
            scratch == nullptr) ? 1 : 0;
  wgrad_geometry(N, OH, OW, Cout, Cin, in_mul_h, in_mul_w, p.pair ? (num_taps + 1) / 2 : num_taps, &p, &bnw);
  p.in_mul_h = in_mul_h; p.in_mul_w = in_mul_w;
  p.Cin = Cin; p.Cout = Cout; p.Ktot = Ktot; p.num_taps = num_taps;
  p.dw = dw;
  p.scratch = scratch;
  p.db = db;
  for (int t = 0; t < num_taps; ++t) {
    p.tap_dh[t] = (int16_t)tap_dh[t]; p.tap_dw[t] = (int16_t)tap_dw[t]; p.tap_w[t] = (int16_t)tap_w[t];
  }
  CUtensorMapSwizzle wgrad_swizzle = CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B;
  {
    // MN-major TF32 operands must use the 128B-swizzle-with-32B-atoms layout (TMA: SWIZZLE_128B_ATOM_32B, UMMA
    // layout type 1): rows of 128 B (32 elements along M/N), 32-byte chunks XOR-ed with (row % 4).  LBO = byte distance
    // between 32-element MN chunks (one TMA box), SBO = byte distance between 4-row K groups, one K=8 MMA = 1024 B.
    uint32_t lbo = kWPix * 128, sbo = 512, a_mn = 1, b_mn = 1, kstep = 1024, layout = 1;
#ifdef DCNN_DEBUG   // layout experiments without a rebuild (scripts/dbg_wgrad.py): never in release builds
    uint32_t swz = 4;
    if (const char* dbg = getenv("DCNN_WGRAD_DBG"))
      sscanf(dbg, "%u,%u,%u,%u,%u,%u,%u", &lbo, &sbo, &a_mn, &b_mn, &kstep, &layout, &swz);
    wgrad_swizzle = (CUtensorMapSwizzle)swz;
#endif
    p.desc_hi_a = make_smem_desc_hi(lbo, sbo, layout);
    p.desc_hi_b = make_smem_desc_hi(lbo, sbo, layout);
    p.idesc = make_idesc_tf32(128, bnw, a_mn, b_mn);
    p.idesc_bias = make_idesc_tf32(128, 16, a_mn, b_mn);
    p.k_step_bytes = kstep;
  }
  CUtensorMap tmDY, tmX;
  int rc = make_tmap_nhwc(&tmDY, dy, N, OH, OW, Cout, p.bn, p.bh, p.bw, 1, 1, wgrad_swizzle);
  if (rc) return rc;
  rc = make_tmap_nhwc(&tmX, x, N, H, W, Cin, p.bn, p.bh, p.bw, in_mul_h, in_mul_w, wgrad_swizzle);
  if (rc) return rc;
  rc = bnw == 128 ? launch_wgrad<128, 3>(tmDY, tmX, p, stream) : launch_wgrad<64, 4>(tmDY, tmX, p, stream);
  if (rc) return rc;
  if (scratch != nullptr) {
    const long long total4 = (long long)Cout * Ktot / 4;   // Ktot % 4 == 0 since Cin % 4 == 0
    wgrad_reduce_kernel<<<ew_grid(total4, 256), 256, 0, stream>>>(scratch, dw, p.splits, total4);
    DCNN_CHECK_LAUNCH();
    DCNN_COUNT_LAUNCH(1);
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// TMA semantics probe (used by the GPU tests): loads ONE 4-D box at the given start coordinate and dumps the raw
// shared-memory image (swizzled, bn*bh*bw rows x 128 B) to `out`.
namespace dcnn {
__global__ void tma_probe_kernel(const __grid_constant__ CUtensorMap tm, int c0, int c1, int c2, int c3, int bytes,
                                 float* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment for the 128B swizzle, computed as an offset so that the compiler keeps the shared address space
  // (a pointer rebuilt from an integer becomes generic: LD.E / ST.E instead of LDS / STS in the epilogue)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar;
  float* f = reinterpret_cast<float*>(smem);
  for (int i = threadIdx.x; i < bytes / 4; i += blockDim.x) f[i] = -12345.f;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  fence_proxy_async();
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bar, bytes);
    tma_load_4d(smem, &tm, &bar, c0, c1, c2, c3);
  }
  mbar_wait(&bar, 0);
  for (int i = threadIdx.x; i < bytes / 4; i += blockDim.x) out[i] = f[i];
}
}  // namespace dcnn

DCNN_API int dcnn_tma_probe(const float* x, int N, int H, int W, int C, int bn, int bh, int bw, int mul_h, int mul_w,
                            int c0, int c1, int c2, int c3, float* out, cudaStream_t stream) {
  CUtensorMap tm;
  int rc = make_tmap_nhwc(&tm, x, N, H, W, C, bn, bh, bw, mul_h, mul_w);
  if (rc) return rc;
  const int bytes = bn * bh * bw * 128;
  DCNN_CHECK_CUDA(cudaFuncSetAttribute(tma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes + 1024));
  tma_probe_kernel<<<1, 128, bytes + 1024, stream>>>(tm, c0, c1, c2, c3, bytes, out);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// TMA ingest probe (measurement tool, scripts/tma_ingest_probe.py): one persistent CTA per SM streams operand
// tiles through the same mbarrier ring as the convolution kernels but issues no MMA -- the consumer just hands
// the stage back.  It answers the question the conv_tc2 capture raised (profiles/03): is the 43 B/clk/SM operand
// feed a per-SM ingest limit or the chip-wide L2 limit, and how much of it disappears when all CTAs read the SAME
// B tile (what a cluster multicast would exploit)?  Each iteration loads an A box {32, a_rows} that is unique to the
// CTA and a B box {32, b_rows} that is either shared by all CTAs (b_shared = 1) or unique as well.
namespace dcnn {
__global__ void __launch_bounds__(64, 1) tma_stream_probe_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                 const __grid_constant__ CUtensorMap tmB, int iters,
                                                                 int stages, int a_rows, int b_rows, int a_row_blocks,
                                                                 int a_col_blocks, int b_row_blocks, int b_col_blocks,
                                                                 int b_shared) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment for the 128B swizzle, computed as an offset so that the compiler keeps the shared address space
  // (a pointer rebuilt from an integer becomes generic: LD.E / ST.E instead of LDS / STS in the epilogue)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t full_bar[8], empty_bar[8];
  const int stage_bytes = (a_rows + b_rows) * 128;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    fence_barrier_init();
  }
  __syncthreads();
  if (warp == 0 && lane == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (int it = 0; it < iters; ++it) {
      // walk memory the way the convolution does: a CTA owns a block of consecutive rows (pixels) and visits its
      // 32-column chunks on consecutive iterations (the k-steps), then moves to its next row block.  (The first version
      // of this probe used 4096-column matrices and jumped to a new 2 MB region every iteration; its timings were
      // dominated by that, see profiles/04.)
      const int kc = it % a_col_blocks;
      const long long tile = (long long)(it / a_col_blocks) * gridDim.x + blockIdx.x;
      const int a_cb = kc, a_rb = (int)(tile % a_row_blocks);
      const int b_cb = it % b_col_blocks;
      const int b_rb = b_shared ? 0 : (int)(tile % b_row_blocks);
      mbar_wait(&empty_bar[stage], phase ^ 1);
      mbar_arrive_expect_tx(&full_bar[stage], stage_bytes);
      uint8_t* s = smem + (size_t)stage * stage_bytes;
      tma_load_2d(s, &tmA, &full_bar[stage], a_cb * 32, a_rb * a_rows);
      if (b_rows > 0) tma_load_2d(s + a_rows * 128, &tmB, &full_bar[stage], b_cb * 32, b_rb * b_rows);
      if (++stage == stages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1 && lane == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (int it = 0; it < iters; ++it) {
      mbar_wait(&full_bar[stage], phase);
      mbar_arrive(&empty_bar[stage]);
      if (++stage == stages) { stage = 0; phase ^= 1; }
    }
  }
}
}  // namespace dcnn

// A: [a_total_rows][a_cols], B: [b_total_rows][b_cols] row-major fp32 (cols % 32 == 0); boxes of a_rows / b_rows rows
// x 32 columns (rows <= 256; b_rows = 0 skips B).  `ctas` <= 0 means one per SM.  Returns the bytes each CTA loaded
// per iteration through *bytes_per_iter (may be NULL).
DCNN_API int dcnn_tma_stream_probe(const float* A, long long a_total_rows, long long a_cols, int a_rows, const float* B,
                                   long long b_total_rows, long long b_cols, int b_rows, int iters, int stages,
                                   int b_shared, int ctas, long long* bytes_per_iter, cudaStream_t stream) {
  DCNN_REQUIRE(a_rows >= 8 && a_rows <= 256 && b_rows >= 0 && b_rows <= 256 && a_rows % 8 == 0 && b_rows % 8 == 0,
               "tma_stream_probe: box rows must be multiples of 8 in [8,256] (a %d, b %d)", a_rows, b_rows);
  DCNN_REQUIRE(stages >= 1 && stages <= 8, "tma_stream_probe: stages %d out of range", stages);
  const int stage_bytes = (a_rows + b_rows) * 128;
  const int smem = stages * stage_bytes + 1024;
  DCNN_REQUIRE(smem <= 227 * 1024, "tma_stream_probe: %d bytes of shared memory requested", smem);
  CUtensorMap tmA, tmB;
  int rc = make_tmap_2d(&tmA, A, a_total_rows, a_cols, a_rows);
  if (rc) return rc;
  rc = make_tmap_2d(&tmB, b_rows > 0 ? B : A, b_rows > 0 ? b_total_rows : a_total_rows, b_rows > 0 ? b_cols : a_cols,
                    b_rows > 0 ? b_rows : a_rows);
  if (rc) return rc;
  DCNN_CHECK_CUDA(cudaFuncSetAttribute(tma_stream_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  if (ctas <= 0) ctas = num_sms();
  tma_stream_probe_kernel<<<ctas, 64, smem, stream>>>(tmA, tmB, iters, stages, a_rows, b_rows,
                                                      (int)(a_total_rows / a_rows), (int)(a_cols / 32),
                                                      b_rows > 0 ? (int)(b_total_rows / b_rows) : 1,
                                                      b_rows > 0 ? (int)(b_cols / 32) : 1, b_shared);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  if (bytes_per_iter) *bytes_per_iter = stage_bytes;
  return 0;
}
