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

// One optimizer step on four consecutive elements: gradient g (already averaged), parameter w, state at s (local shard).
__device__ __forceinline__ float4 ar_update(const ArParams& p, float lr, float bc1, float bc2, float4 g, float4 w,
                                            long long s) {
  float gg[4] = {g.x, g.y, g.z, g.w};
  float ww[4] = {w.x, w.y, w.z, w.w};
  if (p.mode == 0) {
#pragma unroll
    for (int j = 0; j < 4; ++j) ww[j] -= lr * gg[j];
  } else if (p.mode == 1) {
    float4 v = *reinterpret_cast<float4*>(p.s1 + s);
    float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) { vv[j] = p.a * vv[j] - lr * gg[j]; ww[j] += vv[j]; }
    *reinterpret_cast<float4*>(p.s1 + s) = make_float4(vv[0], vv[1], vv[2], vv[3]);
  } else {
    float4 m4 = *reinterpret_cast<float4*>(p.s1 + s), v4 = *reinterpret_cast<float4*>(p.s2 + s);
    float mm[4] = {m4.x, m4.y, m4.z, m4.w}, vv[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      mm[j] = p.a * mm[j] + (1.f - p.a) * gg[j];
      vv[j] = p.b * vv[j] + (1.f - p.b) * gg[j] * gg[j];
      float upd = lr * (mm[j] / bc1) / (sqrtf(vv[j] / bc2) + p.eps);
      if (p.wd > 0.f) {
        if (p.mode == 3) ww[j] -= p.wd * lr * ww[j];
        else upd += p.wd * lr * ww[j];
      }
      ww[j] -= upd;
    }
    *reinterpret_cast<float4*>(p.s1 + s) = make_float4(mm[0], mm[1], mm[2], mm[3]);
    *reinterpret_cast<float4*>(p.s2 + s) = make_float4(vv[0], vv[1], vv[2], vv[3]);
  }
  return make_float4(ww[0], ww[1], ww[2], ww[3]);
}

// W = the number of replicas when it is 2, 4 or 8: a thread handles TWO 16-byte slices per iteration and issues all
// 2 * (W - 1) peer loads before the first one is consumed -- an NVLink load takes microseconds, and a load-add-load-add
// chain would keep only one of them in flight per thread.  W = 0: any number of replicas, plain load-add loop.  The
// summation order (own, rank+1, rank+2, ...) is the same in both, so the results are bit-identical.
// The rendezvous epoch is (*epoch_counter + 1); the last block to finish stores it back, so the counter needs no kernel of
// its own and the launch replays from a CUDA graph.
template <int W>
__global__ void __launch_bounds__(512) fused_allreduce_update_kernel(const __grid_constant__ ArParams pp) {
  const unsigned long long epoch = *pp.epoch_ptr + 1;
  float lr = pp.lr, bc1 = pp.bc1, bc2 = pp.bc2;
  if (pp.hyper != nullptr) { lr = pp.hyper[0]; bc1 = pp.hyper[1]; bc2 = pp.hyper[2]; }
  const int rank = pp.rank, world = pp.world;
  // every block waits until all replicas have finished producing gradients for this epoch
  if (threadIdx.x == 0) {
    if (blockIdx.x == 0) peer_signal(pp.start_flags, rank, world, epoch);
    peer_wait(pp.start_flags, rank, world, epoch);
  }
  __syncthreads();

  const float* gown = pp.grads[rank];
  const float* pown = pp.params[rank];
  const long long n4 = (pp.shard_end - pp.shard_begin) >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const float inv = pp.inv_world;
  if (W != 0) {
    constexpr int WW = W > 0 ? W : 1;
    const float* gpeer[WW];
#pragma unroll
    for (int r = 1; r < WW; ++r) gpeer[r] = pp.grads[(rank + r) & (WW - 1)];
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += 2 * stride) {
      const long long e0 = pp.shard_begin + i * 4;
      const bool two = i + stride < n4;
      const long long e1 = two ? e0 + stride * 4 : e0;
      float4 v0[WW], v1[WW];
      // reduce-scatter: this 16-byte slice of every replica (own copy first, peers over NVLink)
#pragma unroll
      for (int r = 1; r < WW; ++r) { v0[r] = ld_peer4(gpeer[r] + e0); v1[r] = ld_peer4(gpeer[r] + e1); }
      v0[0] = *reinterpret_cast<const float4*>(gown + e0);
      v1[0] = *reinterpret_cast<const float4*>(gown + e1);
      const float4 w0 = *reinterpret_cast<const float4*>(pown + e0), w1 = *reinterpret_cast<const float4*>(pown + e1);
      float4 g0 = v0[0], g1 = v1[0];
#pragma unroll
      for (int r = 1; r < WW; ++r) {
        g0.x += v0[r].x; g0.y += v0[r].y; g0.z += v0[r].z; g0.w += v0[r].w;
        g1.x += v1[r].x; g1.y += v1[r].y; g1.z += v1[r].z; g1.w += v1[r].w;
      }
      g0.x *= inv; g0.y *= inv; g0.z *= inv; g0.w *= inv;
      g1.x *= inv; g1.y *= inv; g1.z *= inv; g1.w *= inv;
      const float4 o0 = ar_update(pp, lr, bc1, bc2, g0, w0, e0 - pp.shard_begin);
      // all-gather: publish the updated slice to every replica (own copy included)
#pragma unroll
      for (int r = 0; r < WW; ++r) *reinterpret_cast<float4*>(pp.params[(rank + r) & (WW - 1)] + e0) = o0;
      if (two) {
        const float4 o1 = ar_update(pp, lr, bc1, bc2, g1, w1, e1 - pp.shard_begin);
#pragma unroll
        for (int r = 0; r < WW; ++r) *reinterpret_cast<float4*>(pp.params[(rank + r) & (WW - 1)] + e1) = o1;
      }
    }
  } else {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += stride) {
      const long long e = pp.shard_begin + i * 4;
      float4 g = *reinterpret_cast<const float4*>(gown + e);
#pragma unroll 1
      for (int r = 1; r < world; ++r) {
        const float4 v = ld_peer4(pp.grads[(rank + r) % world] + e);
        g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
      }
      g.x *= inv; g.y *= inv; g.z *= inv; g.w *= inv;
      const float4 w = *reinterpret_cast<const float4*>(pown + e);
      const float4 out = ar_update(pp, lr, bc1, bc2, g, w, e - pp.shard_begin);
#pragma unroll 1
      for (int r = 0; r < world; ++r) *reinterpret_cast<float4*>(pp.params[(rank + r) % world] + e) = out;
    }
  }
  // Gradient zeroing: a rank may only zero ITS OWN gradient buffer once every peer has read it, i.e. after the "done"
  // signals below have arrived from all of them: wait_done_zero_kernel does that.
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();   // the block's parameter stores (ordered before this by the barrier) reach the peers first
    const unsigned int done = atomicAdd(pp.block_counter, 1u);
    if (done == gridDim.x - 1) {
      *pp.block_counter = 0;
      *const_cast<unsigned long long*>(pp.epoch_ptr) = epoch;
      peer_signal(pp.done_flags, rank, world, epoch);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Push variant of the fused all-reduce + update: the same arithmetic in the same order, but nothing is ever LOADED
// over NVLink.  (Measured on 2 GPUs, profiles/06: the pull kernel needs 156 us against a 124 us NVLink floor because
// every gradient element waits for a peer-load round trip, and the peers may only zero their gradients after a second
// kernel has seen everybody's "done" flag: 51 us more.)
//   phase A  every rank copies the gradient slice that replica o owns into o's staging buffer (posted NVLink stores)
//            and zeroes that slice locally -- no peer ever reads another replica's gradients;
//   barrier  "my slices have arrived" (fence + epoch flags), every CTA waits for all replicas;
//   phase B  own gradient slice + the R-1 staged slices (all LOCAL loads, summed own, rank+1, rank+2, ... as before)
//            -> 1/R -> optimizer update of the owned shard -> new parameters stored into every replica (posted NVLink
//            stores) -> own gradient slice zeroed;
//   done     fence + epoch flags; block 0 leaves only when every replica has signalled, so the kernel's completion means
//            this replica's parameters are complete and its staging buffer may be overwritten by the next step.
// One launch (no counter kernel, no zeroing kernel).  All CTAs must be co-resident (they wait for each other's
// replicas): the grid is at most one CTA per SM.
struct ArPushParams {
  ArParams base;
  float* stage[kMaxPeers];                 // replica r's staging buffer as mapped here
  long long sb[kMaxPeers], se[kMaxPeers];  // shard [sb, se) of replica r inside this launch's range (multiples of 4)
  long long soff[kMaxPeers];               // offset (floats) of this launch's region in replica r's staging buffer
};

template <int W>
__global__ void __launch_bounds__(512) fused_allreduce_update_push_kernel(const __grid_constant__ ArPushParams q) {
  const ArParams& pp = q.base;
  const unsigned long long epoch = *pp.epoch_ptr + 1;
  float lr = pp.lr, bc1 = pp.bc1, bc2 = pp.bc2;
  if (pp.hyper != nullptr) { lr = pp.hyper[0]; bc1 = pp.hyper[1]; bc2 = pp.hyper[2]; }
  const int rank = pp.rank, world = W > 0 ? W : pp.world;
  float* gown = pp.grads[rank];
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);

  // ---- phase A: push the slices the other replicas own, zero them here
  for (int k = 1; k < world; ++k) {
    const int o = (rank + k) % world;
    const long long b = q.sb[o], len = q.se[o] - b, l4 = len >> 2;
    float* dst = q.stage[o] + q.soff[o] + (long long)(world - k - 1) * len;   // slot of source `rank` at owner o
    float* src = gown + b;
    long long i = tid;
    for (; i + 3 * stride < l4; i += 4 * stride) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = *reinterpret_cast<const float4*>(src + (i + u * stride) * 4);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        *reinterpret_cast<float4*>(dst + (i + u * stride) * 4) = v[u];
        *reinterpret_cast<float4*>(src + (i + u * stride) * 4) = zero4;
      }
    }
    for (; i < l4; i += stride) {
      const float4 v = *reinterpret_cast<const float4*>(src + i * 4);
      *reinterpret_cast<float4*>(dst + i * 4) = v;
      *reinterpret_cast<float4*>(src + i * 4) = zero4;
    }
  }
  // ---- every replica's slices have arrived
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    if (atomicAdd(pp.block_counter, 1u) == gridDim.x - 1) {
      *pp.block_counter = 0;
      peer_signal(pp.start_flags, rank, world, epoch);
    }
    peer_wait(pp.start_flags, rank, world, epoch);
  }
  __syncthreads();

  // ---- phase B: reduce + update the owned shard, publish the parameters
  {
    const long long b = q.sb[rank], len = q.se[rank] - b, l4 = len >> 2;
    const float* st = q.stage[rank] + q.soff[rank];
    const float* pown = pp.params[rank];
    const float inv = pp.inv_world;
    for (long long i = tid; i < l4; i += stride) {
      const long long e = b + i * 4;
      float4 g = *reinterpret_cast<const float4*>(gown + e);
      if constexpr (W > 0) {
        float4 v[W];
#pragma unroll
        for (int k = 0; k < W - 1; ++k) v[k] = ld_peer4(st + (long long)k * len + i * 4);
#pragma unroll
        for (int k = 0; k < W - 1; ++k) { g.x += v[k].x; g.y += v[k].y; g.z += v[k].z; g.w += v[k].w; }
      } else {
#pragma unroll 1
        for (int k = 0; k < world - 1; ++k) {
          const float4 v = ld_peer4(st + (long long)k * len + i * 4);
          g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
        }
      }
      g.x *= inv; g.y *= inv; g.z *= inv; g.w *= inv;
      const float4 w = *reinterpret_cast<const float4*>(pown + e);
      const float4 out = ar_update(pp, lr, bc1, bc2, g, w, i * 4);
      if constexpr (W > 0) {
#pragma unroll
        for (int r = 0; r < W; ++r) *reinterpret_cast<float4*>(pp.params[(rank + r) & (W - 1)] + e) = out;
      } else {
#pragma unroll 1
        for (int r = 0; r < world; ++r) *reinterpret_cast<float4*>(pp.params[(rank + r) % world] + e) = out;
      }
      *reinterpret_cast<float4*>(gown + e) = zero4;
    }
  }
  // ---- done
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    if (atomicAdd(pp.block_counter + 1, 1u) == gridDim.x - 1) {
      pp.block_counter[1] = 0;
      *const_cast<unsigned long long*>(pp.epoch_ptr) = epoch;
      peer_signal(pp.done_flags, rank, world, epoch);
    }
    if (blockIdx.x == 0) peer_wait(pp.done_flags, rank, world, epoch);
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

// Push variant (see fused_allreduce_update_push_kernel).  grad_ptrs / param_ptrs point at the start of this launch's
// range in every replica; stage_ptrs[r] = replica r's staging buffer; sb/se/soff[r]: replica r's shard inside the range
// and the offset of this launch's region in r's staging buffer (host arrays of `world` entries); block_counters: two
// zero-initialised u32.  num_blocks is clamped to one CTA per SM.
DCNN_API int dcnn_fused_allreduce_update_push(float* const* grad_ptrs, float* const* param_ptrs, float* const* stage_ptrs,
                                              const long long* sb, const long long* se, const long long* soff,
                                              float* s1, float* s2, long long n,
                                              unsigned long long* const* start_flag_ptrs,
                                              unsigned long long* const* done_flag_ptrs, unsigned int* block_counters,
                                              int rank, int world, unsigned long long* epoch_counter, int mode, float lr,
                                              float a, float b, float eps, float bc1, float bc2, float wd,
                                              const float* hyper, int num_blocks, cudaStream_t stream) {
  ArPushParams q;
  memset(&q, 0, sizeof(q));
  ArParams& p = q.base;
  DCNN_REQUIRE(world >= 1 && world <= kMaxPeers, "world size %d out of range (max %d)", world, kMaxPeers);
  for (int r = 0; r < world; ++r) {
    DCNN_REQUIRE((sb[r] & 3) == 0 && (se[r] & 3) == 0 && (soff[r] & 3) == 0 && sb[r] <= se[r] && se[r] <= n,
                 "fused all-reduce (push): bad shard of rank %d", r);
    p.grads[r] = grad_ptrs[r]; p.params[r] = param_ptrs[r]; q.stage[r] = stage_ptrs[r];
    q.sb[r] = sb[r]; q.se[r] = se[r]; q.soff[r] = soff[r];
  }
  int rc = fill_flags(&p.start_flags, start_flag_ptrs, world);
  if (rc) return rc;
  rc = fill_flags(&p.done_flags, done_flag_ptrs, world);
  if (rc) return rc;
  p.s1 = s1; p.s2 = s2; p.n = n; p.shard_begin = sb[rank]; p.shard_end = se[rank];
  p.block_counter = block_counters; p.rank = rank; p.world = world; p.epoch_ptr = epoch_counter; p.hyper = hyper;
  p.lr = lr; p.a = a; p.b = b; p.eps = eps; p.bc1 = bc1; p.bc2 = bc2; p.wd = wd; p.inv_world = 1.f / (float)world;
  p.mode = mode;
  // every CTA must be resident (they wait for each other's replicas): at most what the occupancy calculator allows
  int per_sm = 1;
  {
    auto kern = world == 2 ? fused_allreduce_update_push_kernel<2>
                           : (world == 4 ? fused_allreduce_update_push_kernel<4>
                                         : (world == 8 ? fused_allreduce_update_push_kernel<8>
                                                       : fused_allreduce_update_push_kernel<0>));
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 512, 0) != cudaSuccess || per_sm < 1) {
      cudaGetLastError();
      per_sm = 1;
    }
  }
  // default: two CTAs per SM for up to four replicas when they fit (0.199 -> 0.183 ms at R = 2: more posted stores in
  // flight per SM); at R = 8 one CTA per SM measured the same or better (0.347 vs 0.359-0.374 ms)
  if (num_blocks <= 0) num_blocks = (per_sm >= 2 && world <= 4 ? 2 : 1) * num_sms();
  if (num_blocks > per_sm * num_sms()) num_blocks = per_sm * num_sms();
  if (world == 2) fused_allreduce_update_push_kernel<2><<<num_blocks, 512, 0, stream>>>(q);
  else if (world == 4) fused_allreduce_update_push_kernel<4><<<num_blocks, 512, 0, stream>>>(q);
  else if (world == 8) fused_allreduce_update_push_kernel<8><<<num_blocks, 512, 0, stream>>>(q);
  else fused_allreduce_update_push_kernel<0><<<num_blocks, 512, 0, stream>>>(q);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
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
  const int w = p.generic_gather ? 0 : world;
  if (w == 2) fused_allreduce_update_kernel<2><<<num_blocks, 512, 0, stream>>>(p);
  else if (w == 4) fused_allreduce_update_kernel<4><<<num_blocks, 512, 0, stream>>>(p);
  else if (w == 8) fused_allreduce_update_kernel<8><<<num_blocks, 512, 0, stream>>>(p);
  else fused_allreduce_update_kernel<0><<<num_blocks, 512, 0, stream>>>(p);
  DCNN_CHECK_LAUNCH();
  // same CTA count as the reduce kernel: a bucket launch on the side stream (32 CTAs) must not fill the SMs that the
  // backward pass is using with threads that spin on the peers' done flags
  wait_done_zero_kernel<<<num_blocks, 512, 0, stream>>>(p.done_flags, rank, world, epoch_counter, grad_ptrs[rank], n);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(2);
  return 0;
}
