// Cross-GPU primitives over NVLink peer mappings: system-scope release/acquire flags with monotonic epochs.
#pragma once
#include "common.cuh"

namespace dcnn {

constexpr int kMaxPeers = 8;

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// peer data must never be served from this SM's L1 (it may hold last step's values): volatile == relaxed.sys
__device__ __forceinline__ float4 ld_peer4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p)
               : "memory");
  return v;
}

struct PeerFlags {
  unsigned long long* flags[kMaxPeers];  // flags[r] = rank r's flag array (kMaxPeers slots), mapped locally
};

// signal: write `epoch` into slot [rank] of every peer's flag array; wait: until every slot of OUR array >= epoch.
__device__ __forceinline__ void peer_signal(const PeerFlags& pf, int rank, int world, unsigned long long epoch) {
  __threadfence_system();
  for (int r = 0; r < world; ++r) st_release_sys(pf.flags[r] + rank, epoch);
}
__device__ __forceinline__ void peer_wait(const PeerFlags& pf, int rank, int world, unsigned long long epoch) {
  for (int r = 0; r < world; ++r) {
    unsigned long long spins = 0;
    while (ld_acquire_sys(pf.flags[rank] + r) < epoch) {
      if (++spins > (1ull << 31)) __trap();  // a lost peer must not hang the GPU forever
    }
  }
}


// A device-resident epoch counter makes every cross-GPU rendezvous replayable from a CUDA graph: a one-thread kernel
// bumps it (stream-ordered) and the kernels that follow read it, so no step-dependent value is baked into launch args.
__global__ void epoch_bump_kernel(unsigned long long* counter);

struct PeerPtrs {
  const double* p[kMaxPeers];
};

}  // namespace dcnn
