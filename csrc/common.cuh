// Shared host/device helpers for libdcnn_b200 (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define DCNN_API extern "C" __attribute__((visibility("default")))

namespace dcnn {

// ------------------------------------------------------------------ error reporting
// Every entry point returns 0 on success or a non-zero code; the text is fetched with dcnn_last_error().
void set_error(const char* fmt, ...);
const char* last_error();

#define DCNN_CHECK_CUDA(expr)                                                              \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      ::dcnn::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return (int)_e ? (int)_e : -1;                                                       \
    }                                                                                      \
  } while (0)

#define DCNN_CHECK_LAUNCH() DCNN_CHECK_CUDA(cudaGetLastError())

#define DCNN_REQUIRE(cond, ...)         \
  do {                                  \
    if (!(cond)) {                      \
      ::dcnn::set_error(__VA_ARGS__);   \
      return -2;                        \
    }                                   \
  } while (0)

// launch counter: number of kernels this library has launched (bench.py's "gpu_launches" claim)
extern unsigned long long g_launch_count;
#define DCNN_COUNT_LAUNCH(n) (::dcnn::g_launch_count += (n))

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute of a kernel: set it once on every device the
// kernel is launched on (a single process may drive one pipeline stage per GPU).  Usage:
//   DCNN_SET_MAX_SMEM_ONCE(bytes, kernel<TEMPLATE, ARGS>);
#define DCNN_SET_MAX_SMEM_ONCE(bytes, ...)                                                                        \
  do {                                                                                                            \
    static bool dcnn_done_[64] = {};                                                                              \
    int dcnn_dev_ = 0;                                                                                            \
    cudaGetDevice(&dcnn_dev_);                                                                                    \
    if (dcnn_dev_ < 0 || dcnn_dev_ >= 64 || !dcnn_done_[dcnn_dev_]) {                                             \
      DCNN_CHECK_CUDA(cudaFuncSetAttribute(__VA_ARGS__, cudaFuncAttributeMaxDynamicSharedMemorySize, (bytes)));   \
      if (dcnn_dev_ >= 0 && dcnn_dev_ < 64) dcnn_done_[dcnn_dev_] = true;                                         \
    }                                                                                                             \
  } while (0)

static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }
static inline int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}
// grid for grid-stride elementwise kernels: a few waves of 148 SMs x resident CTAs
static inline int ew_grid(long long work_items, int threads, int max_ctas_per_sm = 8) {
  long long need = (work_items + threads - 1) / threads;
  long long cap = (long long)num_sms() * max_ctas_per_sm;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

// ------------------------------------------------------------------ device helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
// streaming (evict-first) variants for data touched exactly once
__device__ __forceinline__ float4 ldcs4(const float* p) { return __ldcs(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void stcs4(float* p, float4 v) { __stcs(reinterpret_cast<float4*>(p), v); }

}  // namespace dcnn
