// Data-parallel communication over NVLink 5 / NVSwitch peer memory -- no NCCL on this path.
//
// The reference has no data parallelism at all; this is new capability K26/K27:
//   * symmetric buffers: each rank cudaMalloc's the same layout, exports a CUDA IPC handle, peers map it, so a
//     kernel can dereference every replica's gradients / parameters / BN moments directly;
//   * fused_allreduce_update_kernel: ONE kernel does  reduce-scatter (P2P loads of every replica's gradient
//     slice, summed in registers)  ->  1/R scale  ->  SGD / momentum / Adam / AdamW update of the local shard
//     (optimizer state is sharded: each replica only keeps m,v for the slice it owns)  ->  all-gather (P2P stores of
//     the updated parameters into every replica)  ->  gradient zeroing.  Cross-rank ordering uses monotonic epoch
//     flags written with st.release.sys and polled with ld.acquire.sys;
//   * peer_barrier_kernel: the same flag protocol as a standalone stream-ordered barrier (used by sync-BatchNorm
//     around the exchange of per-channel moments, whose merge runs in norm.cu's *_sync kernels).
#include "peer.cuh"

namespace dcnn {

__global__ void epoch_bump_kernel(unsigned long long* counter) { *counter += 1; }

__global__ void peer_barrier_kernel(PeerFlags pf, int rank, int world, const unsigned long long* epoch_ptr) {
  if (threadIdx.x == 0) {
    const unsigned long long epoch = *epoch_ptr;
    peer_signal(pf, rank, world, epoch);
    peer_wait(pf, rank, world, epoch);
  }
}

struct ArParams {
  float* grads[kMaxPeers];   // every replica's flat gradient buffer (same layout)
  float* params[kMaxPeers];  // every replica's flat parameter buffer
  float* s1;                 // local shard state (velocity / m), indexed from shard_begin
  float* s2;                 // local shard state (v)
  long long n;               // total elements
  long long shard_begin, shard_end;  // [begin,end) owned by this rank (multiple of 4)
  PeerFlags start_flags, done_flags;
  unsigned int* block_counter;       // local, zero-initialised; reset by the last block
  int rank, world;
  const unsigned long long* epoch_ptr;  // device-resident, bumped by epoch_bump_kernel just before this kernel
  const float* hyper;                   // optional device {lr, bc1, bc2} (CUDA-graph replay)
  float lr, a, b, eps, bc1, bc2, wd, inv_world;
  int mode;  // 0 SGD, 1 momentum, 2 Adam, 3 AdamW
  int generic_gather;  // 0 (default): loads-first gather for 2 / 4 / 8 replicas; 1: the load-add loop (other world sizes, DCNN_AR_HOISTED=0)
};

// Own slice plus the W-1 peers' slices.  All peer loads are issued before the first one is consumed: an NVLink load
// takes microseconds, and a load-add-load-add chain would keep only one of them in flight per thread.  The summation
// order (own, rank+1, rank+2, ...) is the same as the generic loop's, so the result is bit-identical to it.
template <int W>
__device__ __forceinline__ float4 gather_sum(const ArParams& p, long long e) {
  float4 v[W];
  v[0] = *reinterpret_cast<const float4*>(p.grads[p.rank] + e);
#pragma unroll
  for (int r = 1; r < W; ++r) v[r] = ld_peer4(p.grads[(p.rank + r) & (W - 1)] + e);
  float4 g = v[0];
#pragma unroll
  for (int r = 1; r < W; ++r) { g.x += v[r].x; g.y += v[r].y; g.z += v[r].z; g.w += v[r].w; }
  return g;
}

__global__ void __launch_bounds__(512) fused_allreduce_update_kernel(const __grid_constant__ ArParams pp) {
  ArParams p = pp;
  const unsigned long long epoch = *p.epoch_ptr;
  if (p.hyper != nullptr) { p.lr = p.hyper[0]; p.bc1 = p.hyper[1]; p.bc2 = p.hyper[2]; }
  // every block waits until all replicas have finished producing gradients for this epoch
  if (threadIdx.x == 0) {
    if (blockIdx.x == 0) peer_signal(p.start_flags, p.rank, p.world, epoch);
    peer_wait(p.start_flags, p.rank, p.world, epoch);
  }
  __syncthreads();

  const long long n4 = (p.shard_end - p.shard_begin) >> 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    const long long e = p.shard_begin + i * 4;
    // reduce-scatter: sum this 16-byte slice over all replicas (own copy first, peers over NVLink)
    float4 g;
    switch (p.generic_gather ? 0 : p.world) {
      case 2: g = gather_sum<2>(pp, e); break;
      case 4: g = gather_sum<4>(pp, e); break;
      case 8: g = gather_sum<8>(pp, e); break;
      default: {
        g = *reinterpret_cast<const float4*>(pp.grads[p.rank] + e);
#pragma unroll 1
        for (int r = 1; r < p.world; ++r) {
          const int src = (p.rank + r) % p.world;
          const float4 v = ld_peer4(pp.grads[src] + e);
          g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
        }
      }
    }
    g.x *= p.inv_world; g.y *= p.inv_world; g.z *= p.inv_world; g.w *= p.inv_world;
    float4 w = *reinterpret_cast<const float4*>(pp.params[p.rank] + e);
    float gg[4] = {g.x, g.y, g.z, g.w};
    float ww[4] = {w.x, w.y, w.z, w.w};
    const long long s = e - p.shard_begin;
    if (p.mode == 0) {
#pragma unroll
      for (int j = 0; j < 4; ++j) ww[j] -= p.lr * gg[j];
    } else if (p.mode == 1) {
      float4 v = *reinterpret_cast<float4*>(p.s1 + s);
      float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) { vv[j] = p.a * vv[j] - p.lr * gg[j]; ww[j] += vv[j]; }
      *reinterpret_cast<float4*>(p.s1 + s) = make_float4(vv[0], vv[1], vv[2], vv[3]);
    } else {
      float4 m4 = *reinterpret_cast<float4*>(p.s1 + s), v4 = *reinterpret_cast<float4*>(p.s2 + s);
      float mm[4] = {m4.x, m4.y, m4.z, m4.w}, vv[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        mm[j] = p.a * mm[j] + (1.f - p.a) * gg[j];
        vv[j] = p.b * vv[j] + (1.f - p.b) * gg[j] * gg[j];
        float upd = p.lr * (mm[j] / p.bc1) / (sqrtf(vv[j] / p.bc2) + p.eps);
        if (p.wd > 0.f) {
          if (p.mode == 3) ww[j] -= p.wd * p.lr * ww[j];
          else upd += p.wd * p.lr * ww[j];
        }
        ww[j] -= upd;
      }
      *reinterpret_cast<float4*>(p.s1 + s) = make_float4(mm[0], mm[1], mm[2], mm[3]);
      *reinterpret_cast<float4*>(p.s2 + s) = make_float4(vv[0], vv[1], vv[2], vv[3]);
    }
    const float4 out = make_float4(ww[0], ww[1], ww[2], ww[3]);
    // all-gather: publish the updated slice to every replica (own copy included)
#pragma unroll 1
    for (int r = 0; r < p.world; ++r) {
      const int dst = (p.rank + r) % p.world;
      *reinterpret_cast<float4*>(pp.params[dst] + e) = out;   // pointer tables are read from the parameter bank
    }
  }
  // this rank may only zero ITS OWN gradient buffer once every peer has read it: that is guaranteed after the
  // done barrier below, so zeroing is done by dcnn_peer_barrier + the next backward's producers (grads are
  // overwritten or zero-filled by zero_after_done_kernel).
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int done = atomicAdd(p.block_counter, 1u);
    if (done == gridDim.x - 1) {
      *p.block_counter = 0;
      peer_signal(p.done_flags, p.rank, p.world, epoch);
    }
  }
}

// Waits for every replica's "done" signal of this epoch, then zeroes the local gradient buffer.
__global__ void __launch_bounds__(512) wait_done_zero_kernel(PeerFlags done_flags, int rank, int world,
                                                             const unsigned long long* epoch_ptr, float* grads,
                                                             long long n) {
  if (threadIdx.x == 0) peer_wait(done_flags, rank, world, *epoch_ptr);
  __syncthreads();
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x)
    *reinterpret_cast<float4*>(grads + i * 4) = make_float4(0, 0, 0, 0);
}

}  // namespace dcnn
using namespace dcnn;

// ----------------------------------------------------------------------------------------------- IPC plumbing
DCNN_API int dcnn_dev_malloc(void** ptr, long long bytes) {
  DCNN_CHECK_CUDA(cudaMalloc(ptr, (size_t)bytes));
  DCNN_CHECK_CUDA(cudaMemset(*ptr, 0, (size_t)bytes));
  return 0;
}
DCNN_API int dcnn_dev_free(void* ptr) {
  DCNN_CHECK_CUDA(cudaFree(ptr));
  return 0;
}
DCNN_API int dcnn_ipc_get_handle(void* ptr, unsigned char* handle64) {
  cudaIpcMemHandle_t h;
  DCNN_CHECK_CUDA(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle64, &h, sizeof(h));
  return 0;
}
DCNN_API int dcnn_ipc_open(const unsigned char* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  DCNN_CHECK_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
DCNN_API int dcnn_ipc_close(void* ptr) {
  DCNN_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}
DCNN_API int dcnn_enable_peer_access(int peer_device) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  DCNN_CHECK_CUDA(e);
  return 0;
}

static int fill_flags(PeerFlags* pf, unsigned long long* const* ptrs, int world) {
  DCNN_REQUIRE(world >= 1 && world <= kMaxPeers, "world size %d out of range (max %d)", world, kMaxPeers);
  for (int r = 0; r < kMaxPeers; ++r) pf->flags[r] = r < world ? ptrs[r] : nullptr;
  return 0;
}

// flag_ptrs[r]: rank r's flag array (>= 8 u64, zero-initialised) as mapped in THIS process
DCNN_API int dcnn_peer_barrier(unsigned long long* const* flag_ptrs, int rank, int world,
                               unsigned long long* epoch_counter, cudaStream_t stream) {
  PeerFlags pf;
  int rc = fill_flags(&pf, flag_ptrs, world);
  if (rc) return rc;
  epoch_bump_kernel<<<1, 1, 0, stream>>>(epoch_counter);
  peer_barrier_kernel<<<1, 32, 0, stream>>>(pf, rank, world, epoch_counter);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(2);
  return 0;
}

DCNN_API int dcnn_fused_allreduce_update(float* const* grad_ptrs, float* const* param_ptrs, float* s1, float* s2,
                                         long long n, long long shard_begin, long long shard_end,
                                         unsigned long long* const* start_flag_ptrs,
                                         unsigned long long* const* done_flag_ptrs, unsigned int* block_counter,
                                         int rank, int world, unsigned long long* epoch_counter, int mode, float lr,
                                         float a, float b, float eps, float bc1, float bc2, float wd,
                                         const float* hyper, int num_blocks, cudaStream_t stream) {
  DCNN_REQUIRE((shard_begin & 3) == 0 && (shard_end & 3) == 0 && (n & 3) == 0,
               "fused all-reduce needs 4-element aligned shards");
  ArParams p;
  memset(&p, 0, sizeof(p));
  for (int r = 0; r < world; ++r) { p.grads[r] = grad_ptrs[r]; p.params[r] = param_ptrs[r]; }
  int rc = fill_flags(&p.start_flags, start_flag_ptrs, world);
  if (rc) return rc;
  rc = fill_flags(&p.done_flags, done_flag_ptrs, world);
  if (rc) return rc;
  p.s1 = s1; p.s2 = s2; p.n = n; p.shard_begin = shard_begin; p.shard_end = shard_end;
  p.block_counter = block_counter; p.rank = rank; p.world = world; p.epoch_ptr = epoch_counter; p.hyper = hyper;
  p.lr = lr; p.a = a; p.b = b; p.eps = eps; p.bc1 = bc1; p.bc2 = bc2; p.wd = wd; p.inv_world = 1.f / (float)world;
  p.mode = mode;
  {
    static const bool hoisted = !(getenv("DCNN_AR_HOISTED") != nullptr && atoi(getenv("DCNN_AR_HOISTED")) == 0);
    p.generic_gather = hoisted ? 0 : 1;
  }
  if (num_blocks <= 0) num_blocks = num_sms();
  epoch_bump_kernel<<<1, 1, 0, stream>>>(epoch_counter);
  fused_allreduce_update_kernel<<<num_blocks, 512, 0, stream>>>(p);
  DCNN_CHECK_LAUNCH();
  // same CTA count as the reduce kernel: a bucket launch on the side stream (32 CTAs) must not fill the SMs that the
  // backward pass is using with threads that spin on the peers' done flags
  wait_done_zero_kernel<<<num_blocks, 512, 0, stream>>>(p.done_flags, rank, world, epoch_counter, grad_ptrs[rank], n);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(3);
  return 0;
}
