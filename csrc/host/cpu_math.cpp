// Native CPU math for the host-side path: blocked AVX2/FMA GEMM (float and double, all four transpose
// combinations), vectorised elementwise primitives, reductions and a 4x-unrolled FMA dot product.
//
// Role: the reference ships hand-written AVX2 kernels for its CPU backend (src/math/cpu/sgemm.cpp, dgemm.cpp;
// src/ops/cpu/skernels.cpp, dkernels.cpp; src/utils/simd_dot_product.S) driven by a TBB/OpenMP parallel_for
// (include/threading/thread_handler.hpp).  This file is the equivalent for this framework's CPU path (the MNIST-CPU
// configuration and the numerical oracle): a BLIS-style GEMM -- pack A into MR-row strips and B into NR-column strips,
// 6x16 (float) / 6x8 (double) register-tiled FMA micro-kernel, KC x MC x NC cache blocking, OpenMP over the MC blocks
// -- rather than the reference's 32^3 blocked loops.  Everything is row-major with explicit leading dimensions.
#include <immintrin.h>
#include <omp.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <vector>

#include "host_common.h"

namespace {

// ------------------------------------------------------------------------------------------------ micro-kernels
template <typename T>
struct Micro;

template <>
struct Micro<float> {
  static constexpr int MR = 6, NR = 16;
  // acc[MR][NR] = sum_k a[k*MR + i] * b[k*NR + j]
  static inline void run(int kc, const float* a, const float* b, float* acc) {
    __m256 c[MR][2];
    for (int i = 0; i < MR; ++i) c[i][0] = c[i][1] = _mm256_setzero_ps();
    for (int k = 0; k < kc; ++k) {
      const __m256 b0 = _mm256_loadu_ps(b), b1 = _mm256_loadu_ps(b + 8);
#pragma GCC unroll 6
      for (int i = 0; i < MR; ++i) {
        const __m256 ai = _mm256_broadcast_ss(a + i);
        c[i][0] = _mm256_fmadd_ps(ai, b0, c[i][0]);
        c[i][1] = _mm256_fmadd_ps(ai, b1, c[i][1]);
      }
      a += MR;
      b += NR;
    }
    for (int i = 0; i < MR; ++i) {
      _mm256_storeu_ps(acc + i * NR, c[i][0]);
      _mm256_storeu_ps(acc + i * NR + 8, c[i][1]);
    }
  }
  // C[i][0..NR) += alpha * (A-strip x B-strip)[i], straight into the output rows
  static inline void run_direct(int kc, const float* a, const float* b, float* C, long long ldc, float alpha) {
    __m256 c[MR][2];
    for (int i = 0; i < MR; ++i) c[i][0] = c[i][1] = _mm256_setzero_ps();
    for (int k = 0; k < kc; ++k) {
      const __m256 b0 = _mm256_loadu_ps(b), b1 = _mm256_loadu_ps(b + 8);
#pragma GCC unroll 6
      for (int i = 0; i < MR; ++i) {
        const __m256 ai = _mm256_broadcast_ss(a + i);
        c[i][0] = _mm256_fmadd_ps(ai, b0, c[i][0]);
        c[i][1] = _mm256_fmadd_ps(ai, b1, c[i][1]);
      }
      a += MR;
      b += NR;
    }
    const __m256 va = _mm256_set1_ps(alpha);
    for (int i = 0; i < MR; ++i) {
      float* r = C + i * ldc;
      _mm256_storeu_ps(r, _mm256_fmadd_ps(va, c[i][0], _mm256_loadu_ps(r)));
      _mm256_storeu_ps(r + 8, _mm256_fmadd_ps(va, c[i][1], _mm256_loadu_ps(r + 8)));
    }
  }
};

// AVX-512 variant (12 x 32 tile: 24 zmm accumulators), selected at run time when the CPU has avx512f.  The functions
// carry their own target attribute so the rest of the library keeps the portable x86-64-v3 baseline.
#define DCNN_AVX512 __attribute__((target("avx512f,fma")))
struct MicroF512 {
  static constexpr int MR = 12, NR = 32;
  DCNN_AVX512 static inline void mac(int kc, const float* a, const float* b, __m512 (&c)[MR][2]) {
    for (int i = 0; i < MR; ++i) c[i][0] = c[i][1] = _mm512_setzero_ps();
    for (int k = 0; k < kc; ++k) {
      const __m512 b0 = _mm512_loadu_ps(b), b1 = _mm512_loadu_ps(b + 16);
#pragma GCC unroll 12
      for (int i = 0; i < MR; ++i) {
        const __m512 ai = _mm512_set1_ps(a[i]);
        c[i][0] = _mm512_fmadd_ps(ai, b0, c[i][0]);
        c[i][1] = _mm512_fmadd_ps(ai, b1, c[i][1]);
      }
      a += MR;
      b += NR;
    }
  }
  DCNN_AVX512 static void run(int kc, const float* a, const float* b, float* acc) {
    __m512 c[MR][2];
    mac(kc, a, b, c);
    for (int i = 0; i < MR; ++i) {
      _mm512_storeu_ps(acc + i * NR, c[i][0]);
      _mm512_storeu_ps(acc + i * NR + 16, c[i][1]);
    }
  }
  DCNN_AVX512 static void run_direct(int kc, const float* a, const float* b, float* C, long long ldc, float alpha) {
    __m512 c[MR][2];
    mac(kc, a, b, c);
    const __m512 va = _mm512_set1_ps(alpha);
    for (int i = 0; i < MR; ++i) {
      float* r = C + i * ldc;
      _mm512_storeu_ps(r, _mm512_fmadd_ps(va, c[i][0], _mm512_loadu_ps(r)));
      _mm512_storeu_ps(r + 16, _mm512_fmadd_ps(va, c[i][1], _mm512_loadu_ps(r + 16)));
    }
  }
};

static bool has_avx512() {
  static const bool v = __builtin_cpu_supports("avx512f") && getenv("DCNN_CPU_NO_AVX512") == nullptr;
  return v;
}

template <>
struct Micro<double> {
  static constexpr int MR = 6, NR = 8;
  static inline void run(int kc, const double* a, const double* b, double* acc) {
    __m256d c[MR][2];
    for (int i = 0; i < MR; ++i) c[i][0] = c[i][1] = _mm256_setzero_pd();
    for (int k = 0; k < kc; ++k) {
      const __m256d b0 = _mm256_loadu_pd(b), b1 = _mm256_loadu_pd(b + 4);
#pragma GCC unroll 6
      for (int i = 0; i < MR; ++i) {
        const __m256d ai = _mm256_broadcast_sd(a + i);
        c[i][0] = _mm256_fmadd_pd(ai, b0, c[i][0]);
        c[i][1] = _mm256_fmadd_pd(ai, b1, c[i][1]);
      }
      a += MR;
      b += NR;
    }
    for (int i = 0; i < MR; ++i) {
      _mm256_storeu_pd(acc + i * NR, c[i][0]);
      _mm256_storeu_pd(acc + i * NR + 4, c[i][1]);
    }
  }
  static inline void run_direct(int kc, const double* a, const double* b, double* C, long long ldc, double alpha) {
    alignas(32) double acc[MR * NR];
    run(kc, a, b, acc);
    for (int i = 0; i < MR; ++i)
      for (int j = 0; j < NR; ++j) C[i * ldc + j] += alpha * acc[i * NR + j];
  }
};

// ------------------------------------------------------------------------------------------------ packing
// op(A)(i,k) = A[i*rs + k*cs]; strips of MR rows, zero padded: dst[strip][k][MR]
template <typename T, int MR>
void pack_a(const T* A, long long rs, long long cs, int mc, int kc, T* dst) {
  for (int i0 = 0; i0 < mc; i0 += MR) {
    const int mr = std::min(MR, mc - i0);
    for (int k = 0; k < kc; ++k) {
      for (int i = 0; i < mr; ++i) dst[i] = A[(long long)(i0 + i) * rs + (long long)k * cs];
      for (int i = mr; i < MR; ++i) dst[i] = T(0);
      dst += MR;
    }
  }
}

// op(B)(k,j) = B[k*rs + j*cs]; one strip of NR columns: dst[k][NR]
template <typename T, int NR>
void pack_b_strip(const T* B, long long rs, long long cs, int kc, int nr, T* dst) {
  if (cs == 1 && nr == NR) {
    for (int k = 0; k < kc; ++k) memcpy(dst + (long long)k * NR, B + (long long)k * rs, NR * sizeof(T));
    return;
  }
  for (int k = 0; k < kc; ++k) {
    for (int j = 0; j < nr; ++j) dst[j] = B[(long long)k * rs + (long long)j * cs];
    for (int j = nr; j < NR; ++j) dst[j] = T(0);
    dst += NR;
  }
}

template <typename KERNEL, typename T>
void gemm_impl(int transA, int transB, int M, int N, int K, T alpha, const T* A, int lda, const T* B, int ldb, T beta,
               T* C, int ldc) {
  constexpr int MR = KERNEL::MR, NR = KERNEL::NR;
  constexpr int MC = 96, KC = 256, NC = 2048;   // MC is a multiple of both MR = 6 and MR = 12
  if (M <= 0 || N <= 0) return;
  // C *= beta once, up front (beta == 0 overwrites: NaNs in C must not propagate)
#pragma omp parallel for schedule(static) if ((long long)M * N > 16384)
  for (int i = 0; i < M; ++i) {
    T* c = C + (long long)i * ldc;
    if (beta == T(0)) for (int j = 0; j < N; ++j) c[j] = T(0);
    else if (beta != T(1)) for (int j = 0; j < N; ++j) c[j] *= beta;
  }
  if (K <= 0 || alpha == T(0)) return;
  const long long a_rs = transA ? 1 : lda, a_cs = transA ? lda : 1;
  const long long b_rs = transB ? 1 : ldb, b_cs = transB ? ldb : 1;
  // packing buffers live for the life of the thread: a convolution issues thousands of small products and must not pay
  // an allocation plus a 2 MB clear for each
  static thread_local std::vector<T> bpack_tls;
  const size_t bneed = (size_t)std::min(KC, K) * ((size_t)std::min(NC, N) + NR);
  if (bpack_tls.size() < bneed) bpack_tls.resize(bneed);
  T* const bpack_data = bpack_tls.data();
  const bool par = (2.0 * M * N * K) > 2e6;
#pragma omp parallel if (par)
  {
    static thread_local std::vector<T> apack;
    if (apack.size() < (size_t)(MC + MR) * KC) apack.resize((size_t)(MC + MR) * KC);
    alignas(64) T acc[MR * NR];
    for (int jc = 0; jc < N; jc += NC) {
      const int nc = std::min(NC, N - jc);
      const int strips = (nc + NR - 1) / NR;
      for (int pc = 0; pc < K; pc += KC) {
        const int kc = std::min(KC, K - pc);
#pragma omp for schedule(static)
        for (int s = 0; s < strips; ++s) {
          const int j0 = jc + s * NR;
          pack_b_strip<T, NR>(B + (long long)pc * b_rs + (long long)j0 * b_cs, b_rs, b_cs, kc, std::min(NR, N - j0),
                       bpack_data + (size_t)s * kc * NR);
        }  // implicit barrier: the panel is complete before anyone multiplies
        // tasks = (MC block) x (group of B strips): when M is small the strip groups supply the parallelism (each
        // task re-packs its A block, which is cheap next to the kc * group-width multiply it feeds)
        const int iblocks = (M + MC - 1) / MC;
        int groups = (2 * omp_get_num_threads() + iblocks - 1) / iblocks;
        groups = std::max(1, std::min(groups, strips));
        const int per_group = (strips + groups - 1) / groups;
#pragma omp for collapse(2) schedule(dynamic, 1)
        for (int ib = 0; ib < iblocks; ++ib) {
          for (int g = 0; g < groups; ++g) {
            const int ic = ib * MC, mc = std::min(MC, M - ic);
            const int s_lo = g * per_group, s_hi = std::min(strips, s_lo + per_group);
            if (s_lo >= s_hi) continue;
            pack_a<T, MR>(A + (long long)ic * a_rs + (long long)pc * a_cs, a_rs, a_cs, mc, kc, apack.data());
            for (int s = s_lo; s < s_hi; ++s) {
              const int j0 = jc + s * NR, nr = std::min(NR, N - j0);
              const T* bp = bpack_data + (size_t)s * kc * NR;
              for (int i0 = 0; i0 < mc; i0 += MR) {
                const int mr = std::min(MR, mc - i0);
                T* c = C + (long long)(ic + i0) * ldc + j0;
                if (mr == MR && nr == NR) {   // full tile: accumulate straight into C
                  KERNEL::run_direct(kc, apack.data() + (size_t)(i0 / MR) * kc * MR, bp, c, ldc, alpha);
                } else {
                  KERNEL::run(kc, apack.data() + (size_t)(i0 / MR) * kc * MR, bp, acc);
                  for (int i = 0; i < mr; ++i)
                    for (int j = 0; j < nr; ++j) c[(long long)i * ldc + j] += alpha * acc[i * NR + j];
                }
              }
            }
          }
        }  // implicit barrier: bpack may be overwritten by the next panel
      }
    }
  }
}

template <typename T>
T dot_impl(const T* a, const T* b, long long n);

template <>
float dot_impl<float>(const float* a, const float* b, long long n) {
  __m256 s0 = _mm256_setzero_ps(), s1 = s0, s2 = s0, s3 = s0;
  long long i = 0;
  for (; i + 32 <= n; i += 32) {
    s0 = _mm256_fmadd_ps(_mm256_loadu_ps(a + i), _mm256_loadu_ps(b + i), s0);
    s1 = _mm256_fmadd_ps(_mm256_loadu_ps(a + i + 8), _mm256_loadu_ps(b + i + 8), s1);
    s2 = _mm256_fmadd_ps(_mm256_loadu_ps(a + i + 16), _mm256_loadu_ps(b + i + 16), s2);
    s3 = _mm256_fmadd_ps(_mm256_loadu_ps(a + i + 24), _mm256_loadu_ps(b + i + 24), s3);
  }
  for (; i + 8 <= n; i += 8) s0 = _mm256_fmadd_ps(_mm256_loadu_ps(a + i), _mm256_loadu_ps(b + i), s0);
  s0 = _mm256_add_ps(_mm256_add_ps(s0, s1), _mm256_add_ps(s2, s3));
  alignas(32) float t[8];
  _mm256_store_ps(t, s0);
  float r = ((t[0] + t[1]) + (t[2] + t[3])) + ((t[4] + t[5]) + (t[6] + t[7]));
  for (; i < n; ++i) r += a[i] * b[i];
  return r;
}

template <>
double dot_impl<double>(const double* a, const double* b, long long n) {
  __m256d s0 = _mm256_setzero_pd(), s1 = s0, s2 = s0, s3 = s0;
  long long i = 0;
  for (; i + 16 <= n; i += 16) {
    s0 = _mm256_fmadd_pd(_mm256_loadu_pd(a + i), _mm256_loadu_pd(b + i), s0);
    s1 = _mm256_fmadd_pd(_mm256_loadu_pd(a + i + 4), _mm256_loadu_pd(b + i + 4), s1);
    s2 = _mm256_fmadd_pd(_mm256_loadu_pd(a + i + 8), _mm256_loadu_pd(b + i + 8), s2);
    s3 = _mm256_fmadd_pd(_mm256_loadu_pd(a + i + 12), _mm256_loadu_pd(b + i + 12), s3);
  }
  s0 = _mm256_add_pd(_mm256_add_pd(s0, s1), _mm256_add_pd(s2, s3));
  alignas(32) double t[4];
  _mm256_store_pd(t, s0);
  double r = (t[0] + t[1]) + (t[2] + t[3]);
  for (; i < n; ++i) r += a[i] * b[i];
  return r;
}

// C[M,N] = alpha * A[M,K] * B[N,K]^T + beta * C for SMALL M and N and a LONG K (both operands contiguous along K): the
// weight gradient of a convolution with few channels is exactly this (8 x 25 outputs, K = N*OH*OW = 36 864 for the first
// MNIST layer).  The blocked kernel would spend its time packing and synchronising per 256-deep panel; here every
// thread takes a slice of K, forms its M x N partial with the FMA dot product, and the partials are summed in thread
// order (deterministic for a given thread count).
template <typename T>
void gemm_nt_long_k(int M, int N, int K, T alpha, const T* A, int lda, const T* B, int ldb, T beta, T* C, int ldc) {
  constexpr int kChunk = 2048;
  const int chunks = (K + kChunk - 1) / kChunk;
  const int nt = std::max(1, std::min(omp_get_max_threads(), chunks));
  std::vector<T> part((size_t)nt * M * N, T(0));
#pragma omp parallel num_threads(nt)
  {
    T* mine = part.data() + (size_t)omp_get_thread_num() * M * N;
#pragma omp for schedule(static)
    for (int c = 0; c < chunks; ++c) {
      const int k0 = c * kChunk, len = std::min(kChunk, K - k0);
      for (int i = 0; i < M; ++i)
        for (int j = 0; j < N; ++j)
          mine[i * N + j] += dot_impl<T>(A + (long long)i * lda + k0, B + (long long)j * ldb + k0, len);
    }
  }
  for (int i = 0; i < M; ++i)
    for (int j = 0; j < N; ++j) {
      T s = T(0);
      for (int t = 0; t < nt; ++t) s += part[(size_t)t * M * N + i * N + j];
      T& c = C[(long long)i * ldc + j];
      c = (beta == T(0) ? T(0) : beta * c) + alpha * s;
    }
}

// Parallel reduction of per-chunk results in a fixed order (deterministic for a given thread count-independent
// chunking: the chunk size does not depend on the number of threads).
template <typename T, typename F>
T chunked_sum(long long n, F&& chunk_fn) {
  constexpr long long kChunk = 1 << 16;
  const long long chunks = (n + kChunk - 1) / kChunk;
  if (chunks <= 1) return chunk_fn(0, n);
  std::vector<T> part((size_t)chunks);
#pragma omp parallel for schedule(static)
  for (long long c = 0; c < chunks; ++c) part[(size_t)c] = chunk_fn(c * kChunk, std::min(n, (c + 1) * kChunk));
  T r = T(0);
  for (long long c = 0; c < chunks; ++c) r += part[(size_t)c];
  return r;
}

// op codes shared with dcnn_b200/ops/cpu_native.py
enum BinOp { kAdd = 0, kSub = 1, kMul = 2, kDiv = 3, kMax = 4, kMin = 5 };
enum UnOp { kNeg = 0, kAbs = 1, kSqrt = 2, kRsqrt = 3, kRcp = 4, kExp = 5, kLog = 6, kSquare = 7, kRelu = 8 };

template <typename T>
void binary_impl(int op, const T* a, const T* b, T* y, long long n) {
#define DCNN_BIN(EXPR)                                                          \
  _Pragma("omp parallel for simd schedule(static) if (n > 32768)") for (long long i = 0; i < n; ++i) y[i] = (EXPR); \
  break;
  switch (op) {
    case kAdd: DCNN_BIN(a[i] + b[i])
    case kSub: DCNN_BIN(a[i] - b[i])
    case kMul: DCNN_BIN(a[i] * b[i])
    case kDiv: DCNN_BIN(a[i] / b[i])
    case kMax: DCNN_BIN(a[i] > b[i] ? a[i] : b[i])
    case kMin: DCNN_BIN(a[i] < b[i] ? a[i] : b[i])
    default: break;
  }
#undef DCNN_BIN
}

template <typename T>
void binary_scalar_impl(int op, const T* a, T s, T* y, long long n) {
#define DCNN_BIN(EXPR)                                                          \
  _Pragma("omp parallel for simd schedule(static) if (n > 32768)") for (long long i = 0; i < n; ++i) y[i] = (EXPR); \
  break;
  switch (op) {
    case kAdd: DCNN_BIN(a[i] + s)
    case kSub: DCNN_BIN(a[i] - s)
    case kMul: DCNN_BIN(a[i] * s)
    case kDiv: DCNN_BIN(a[i] / s)
    case kMax: DCNN_BIN(a[i] > s ? a[i] : s)
    case kMin: DCNN_BIN(a[i] < s ? a[i] : s)
    default: break;
  }
#undef DCNN_BIN
}

template <typename T>
void unary_impl(int op, const T* a, T* y, long long n) {
#define DCNN_UN(EXPR)                                                           \
  _Pragma("omp parallel for simd schedule(static) if (n > 32768)") for (long long i = 0; i < n; ++i) y[i] = (EXPR); \
  break;
  switch (op) {
    case kNeg: DCNN_UN(-a[i])
    case kAbs: DCNN_UN(a[i] < T(0) ? -a[i] : a[i])
    case kSqrt: DCNN_UN(std::sqrt(a[i]))
    case kRsqrt: DCNN_UN(T(1) / std::sqrt(a[i]))
    case kRcp: DCNN_UN(T(1) / a[i])
    case kExp: DCNN_UN(std::exp(a[i]))
    case kLog: DCNN_UN(std::log(a[i]))
    case kSquare: DCNN_UN(a[i] * a[i])
    case kRelu: DCNN_UN(a[i] > T(0) ? a[i] : T(0))
    default: break;
  }
#undef DCNN_UN
}

// im2col into the reference's column layout [C*kh*kw][N*OH*OW] (include/tensor/cpu/tensor_ops.hpp:22-170)
void im2col_impl(const float* x, float* col, int N, int C, int H, int W, int kh, int kw, int sh, int sw, int ph, int pw,
                 int OH, int OW) {
  const long long L = (long long)N * OH * OW;
#pragma omp parallel for collapse(2) schedule(static)
  for (int ck = 0; ck < C * kh * kw; ++ck) {
    for (int n = 0; n < N; ++n) {
      const int c = ck / (kh * kw), r = (ck / kw) % kh, s = ck % kw;
      float* dst = col + (long long)ck * L + (long long)n * OH * OW;
      const float* src = x + ((long long)n * C + c) * H * W;
      for (int oh = 0; oh < OH; ++oh) {
        const int ih = oh * sh - ph + r;
        if (ih < 0 || ih >= H) {
          for (int ow = 0; ow < OW; ++ow) dst[oh * OW + ow] = 0.f;
          continue;
        }
        for (int ow = 0; ow < OW; ++ow) {
          const int iw = ow * sw - pw + s;
          dst[oh * OW + ow] = (iw >= 0 && iw < W) ? src[(long long)ih * W + iw] : 0.f;
        }
      }
    }
  }
}

// col2im: gather form (each input element sums the column entries that reference it) so no atomics are needed
void col2im_impl(const float* col, float* x, int N, int C, int H, int W, int kh, int kw, int sh, int sw, int ph, int pw,
                 int OH, int OW) {
  const long long L = (long long)N * OH * OW;
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; ++n) {
    for (int c = 0; c < C; ++c) {
      float* dst = x + ((long long)n * C + c) * H * W;
      for (int h = 0; h < H; ++h) {
        for (int w = 0; w < W; ++w) {
          float acc = 0.f;
          for (int r = 0; r < kh; ++r) {
            const int t = h + ph - r;
            if (t < 0 || t % sh != 0) continue;
            const int oh = t / sh;
            if (oh >= OH) continue;
            for (int s = 0; s < kw; ++s) {
              const int u = w + pw - s;
              if (u < 0 || u % sw != 0) continue;
              const int ow = u / sw;
              if (ow >= OW) continue;
              acc += col[(long long)((c * kh + r) * kw + s) * L + (long long)n * OH * OW + oh * OW + ow];
            }
          }
          dst[(long long)h * W + w] = acc;
        }
      }
    }
  }
}

}  // namespace

// ================================================================================================ C ABI
DCNN_HOST_API void dcnn_cpu_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
DCNN_HOST_API int dcnn_cpu_get_threads() { return omp_get_max_threads(); }

// Row-major GEMM: C[M,N] = alpha * op(A) * op(B) + beta * C.  lda/ldb/ldc are the row strides of the STORED matrices
// (A is stored K x M when transA, B is stored N x K when transB).
DCNN_HOST_API void dcnn_cpu_sgemm(int transA, int transB, int M, int N, int K, float alpha, const float* A, int lda,
                                  const float* B, int ldb, float beta, float* C, int ldc) {
  if (!transA && transB && M <= 64 && N <= 512 && (long long)M * N <= 4096 && K >= 4096 && alpha != 0.f) {
    gemm_nt_long_k<float>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
    return;
  }
  // the 12 x 32 tile only pays off when the problem has enough full tiles; skinny products (a conv with 8 output
  // channels is an 8 x N GEMM) stay on the 6 x 16 kernel
  if (has_avx512() && M >= 48 && N >= 64) gemm_impl<MicroF512, float>(transA, transB, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
  else gemm_impl<Micro<float>, float>(transA, transB, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
}
DCNN_HOST_API void dcnn_cpu_dgemm(int transA, int transB, int M, int N, int K, double alpha, const double* A, int lda,
                                  const double* B, int ldb, double beta, double* C, int ldc) {
  gemm_impl<Micro<double>, double>(transA, transB, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
}

DCNN_HOST_API float dcnn_cpu_sdot(const float* a, const float* b, long long n) {
  return chunked_sum<float>(n, [&](long long lo, long long hi) { return dot_impl<float>(a + lo, b + lo, hi - lo); });
}
DCNN_HOST_API double dcnn_cpu_ddot(const double* a, const double* b, long long n) {
  return chunked_sum<double>(n, [&](long long lo, long long hi) { return dot_impl<double>(a + lo, b + lo, hi - lo); });
}
DCNN_HOST_API double dcnn_cpu_ssum(const float* a, long long n) {
  return chunked_sum<double>(n, [&](long long lo, long long hi) {
    double s = 0;
#pragma omp simd reduction(+ : s)
    for (long long i = lo; i < hi; ++i) s += a[i];
    return s;
  });
}
DCNN_HOST_API double dcnn_cpu_ssum_squared_diff(const float* a, float mean, long long n) {
  return chunked_sum<double>(n, [&](long long lo, long long hi) {
    double s = 0;
#pragma omp simd reduction(+ : s)
    for (long long i = lo; i < hi; ++i) s += (double)(a[i] - mean) * (double)(a[i] - mean);
    return s;
  });
}

DCNN_HOST_API void dcnn_cpu_sbinary(int op, const float* a, const float* b, float* y, long long n) {
  binary_impl<float>(op, a, b, y, n);
}
DCNN_HOST_API void dcnn_cpu_dbinary(int op, const double* a, const double* b, double* y, long long n) {
  binary_impl<double>(op, a, b, y, n);
}
DCNN_HOST_API void dcnn_cpu_sbinary_scalar(int op, const float* a, float s, float* y, long long n) {
  binary_scalar_impl<float>(op, a, s, y, n);
}
DCNN_HOST_API void dcnn_cpu_dbinary_scalar(int op, const double* a, double s, double* y, long long n) {
  binary_scalar_impl<double>(op, a, s, y, n);
}
DCNN_HOST_API void dcnn_cpu_sunary(int op, const float* a, float* y, long long n) { unary_impl<float>(op, a, y, n); }
DCNN_HOST_API void dcnn_cpu_dunary(int op, const double* a, double* y, long long n) { unary_impl<double>(op, a, y, n); }

// y = a*b + c, y = a*b - c, y = c - a*b  (mode 0/1/2)
DCNN_HOST_API void dcnn_cpu_sfma(int mode, const float* a, const float* b, const float* c, float* y, long long n) {
  if (mode == 0) {
#pragma omp parallel for simd schedule(static) if (n > 32768)
    for (long long i = 0; i < n; ++i) y[i] = a[i] * b[i] + c[i];
  } else if (mode == 1) {
#pragma omp parallel for simd schedule(static) if (n > 32768)
    for (long long i = 0; i < n; ++i) y[i] = a[i] * b[i] - c[i];
  } else {
#pragma omp parallel for simd schedule(static) if (n > 32768)
    for (long long i = 0; i < n; ++i) y[i] = c[i] - a[i] * b[i];
  }
}
DCNN_HOST_API void dcnn_cpu_saxpy(float alpha, const float* x, float* y, long long n) {
#pragma omp parallel for simd schedule(static) if (n > 32768)
  for (long long i = 0; i < n; ++i) y[i] += alpha * x[i];
}
DCNN_HOST_API void dcnn_cpu_sset(float* y, float v, long long n) {
#pragma omp parallel for simd schedule(static) if (n > 32768)
  for (long long i = 0; i < n; ++i) y[i] = v;
}

// out[cols][rows] = in[rows][cols], 32x32 blocked
DCNN_HOST_API void dcnn_cpu_stranspose(const float* in, float* out, int rows, int cols) {
  constexpr int B = 32;
#pragma omp parallel for collapse(2) schedule(static) if ((long long)rows * cols > 65536)
  for (int r0 = 0; r0 < rows; r0 += B)
    for (int c0 = 0; c0 < cols; c0 += B)
      for (int r = r0; r < std::min(rows, r0 + B); ++r)
        for (int c = c0; c < std::min(cols, c0 + B); ++c) out[(long long)c * rows + r] = in[(long long)r * cols + c];
}

DCNN_HOST_API void dcnn_cpu_im2col(const float* x, float* col, int N, int C, int H, int W, int kh, int kw, int sh, int sw,
                                   int ph, int pw, int OH, int OW) {
  im2col_impl(x, col, N, C, H, W, kh, kw, sh, sw, ph, pw, OH, OW);
}
DCNN_HOST_API void dcnn_cpu_col2im(const float* col, float* x, int N, int C, int H, int W, int kh, int kw, int sh, int sw,
                                   int ph, int pw, int OH, int OW) {
  col2im_impl(col, x, N, C, H, W, kh, kw, sh, sw, ph, pw, OH, OW);
}
