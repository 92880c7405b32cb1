// NVLS (NVLink SHARP) path of the data-parallel update: gradients and parameters live in memory that is bound to a
// MULTICAST object spanning all replicas, so that
//   * multimem.ld_reduce.add on the multicast address returns the SUM of every replica's gradient element -- the
//     reduction happens inside the NVSwitch, each GPU pulls 1/R of the gradient bytes once instead of (R-1)/R of them
//     from R-1 peers, and
//   * multimem.st on the multicast address writes the updated parameter into every replica with one store.
// The optimizer arithmetic, the sharding of the optimizer state and the epoch-flag rendezvous are those of
// fused_allreduce_update_kernel (comm.cu); only the two transfers change.
//
// Host side: CUDA virtual-memory-management + multicast driver API (resolved through cudaGetDriverEntryPoint, the
// library does not link libcuda).  Rank 0 creates the multicast object and exports it as a POSIX file descriptor; the
// Python layer (parallel/dp.py: NvlsBuffer) passes the descriptor to the other ranks over a UNIX socket.
//
// Opt-in (DCNN_DP_NVLS=1).  Written after the last GPU minutes of its session: NOT yet run on hardware.
#include <cuda.h>
#include <string.h>
#include <unistd.h>

#include "peer.cuh"

namespace dcnn {

// ---------------------------------------------------------------------------------------------- driver entry points
struct DriverApi {
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                               unsigned long long);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
                                         unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*DeviceGet)(CUdevice*, int);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  bool ok;
};

template <typename F>
static bool resolve(F* fn, const char* name) {
  void* ptr = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &ptr, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
      ptr == nullptr) {
    cudaGetLastError();
    return false;
  }
  *fn = reinterpret_cast<F>(ptr);
  return true;
}

static const DriverApi& driver() {
  static DriverApi d = [] {
    DriverApi a;
    memset(&a, 0, sizeof(a));
    a.ok = resolve(&a.MulticastCreate, "cuMulticastCreate") && resolve(&a.MulticastAddDevice, "cuMulticastAddDevice") &&
           resolve(&a.MulticastBindMem, "cuMulticastBindMem") &&
           resolve(&a.MulticastGetGranularity, "cuMulticastGetGranularity") && resolve(&a.MemCreate, "cuMemCreate") &&
           resolve(&a.MemRelease, "cuMemRelease") && resolve(&a.MemAddressReserve, "cuMemAddressReserve") &&
           resolve(&a.MemAddressFree, "cuMemAddressFree") && resolve(&a.MemMap, "cuMemMap") &&
           resolve(&a.MemUnmap, "cuMemUnmap") && resolve(&a.MemSetAccess, "cuMemSetAccess") &&
           resolve(&a.MemExportToShareableHandle, "cuMemExportToShareableHandle") &&
           resolve(&a.MemImportFromShareableHandle, "cuMemImportFromShareableHandle") &&
           resolve(&a.MemGetAllocationGranularity, "cuMemGetAllocationGranularity") &&
           resolve(&a.DeviceGet, "cuDeviceGet") && resolve(&a.DeviceGetAttribute, "cuDeviceGetAttribute");
    return a;
  }();
  return d;
}

#define DCNN_CHECK_CU(expr)                                                         \
  do {                                                                              \
    CUresult r_ = (expr);                                                           \
    DCNN_REQUIRE(r_ == CUDA_SUCCESS, "%s failed with CUresult %d", #expr, (int)r_); \
  } while (0)

}  // namespace dcnn

namespace dcnn {

// ---------------------------------------------------------------------------------------------- device side
__device__ __forceinline__ float4 multimem_ld_reduce_add4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

struct NvlsParams {
  const float* mc_grads;  // multicast view of the flat gradient buffers (same offset in every replica)
  float* mc_params;       // multicast view of the flat parameter buffers
  const float* params;    // this replica's parameters (unicast)
  float* s1;
  float* s2;
  long long shard_begin, shard_end;
  PeerFlags start_flags, done_flags;
  unsigned int* block_counter;
  int rank, world;
  const unsigned long long* epoch_ptr;
  const float* hyper;
  float lr, a, b, eps, bc1, bc2, wd, inv_world;
  int mode;  // 0 SGD, 1 momentum, 2 Adam, 3 AdamW
};

__global__ void __launch_bounds__(512) fused_allreduce_update_nvls_kernel(const __grid_constant__ NvlsParams pp) {
  NvlsParams p = pp;
  const unsigned long long epoch = *p.epoch_ptr;
  if (p.hyper != nullptr) { p.lr = p.hyper[0]; p.bc1 = p.hyper[1]; p.bc2 = p.hyper[2]; }
  if (threadIdx.x == 0) {
    if (blockIdx.x == 0) peer_signal(p.start_flags, p.rank, p.world, epoch);
    peer_wait(p.start_flags, p.rank, p.world, epoch);   // every replica's gradients are complete
  }
  __syncthreads();

  const long long n4 = (p.shard_end - p.shard_begin) >> 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    const long long e = p.shard_begin + i * 4;
    float4 g = multimem_ld_reduce_add4(p.mc_grads + e);   // the switch sums the R replicas' values
    float gg[4] = {g.x * p.inv_world, g.y * p.inv_world, g.z * p.inv_world, g.w * p.inv_world};
    const float4 w = *reinterpret_cast<const float4*>(p.params + e);
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
    multimem_st4(p.mc_params + e, make_float4(ww[0], ww[1], ww[2], ww[3]));   // one store reaches every replica
  }
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

// Waits for every replica's "done" signal of this epoch (all of them have pulled their sums), then zeroes the local
// gradient buffer through its unicast mapping.
__global__ void __launch_bounds__(512) nvls_wait_done_zero_kernel(PeerFlags done_flags, int rank, int world,
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

// ---------------------------------------------------------------------------------------------- host API
// 1 if the device supports multicast objects (NVSwitch fabric + driver), 0 if not, < 0 on error.
DCNN_API int dcnn_nvls_supported(int device) {
  const DriverApi& d = driver();
  if (!d.ok) return 0;
  CUdevice dev;
  if (d.DeviceGet(&dev, device) != CUDA_SUCCESS) return 0;
  int v = 0;
  if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return 0;
  return v ? 1 : 0;
}

static CUmulticastObjectProp mc_prop(int world, size_t bytes) {
  CUmulticastObjectProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.numDevices = (unsigned)world;
  prop.size = bytes;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

// Size (>= bytes) that satisfies both the multicast and the allocation granularity for `world` devices.
DCNN_API long long dcnn_nvls_padded_size(int world, long long bytes, int device) {
  const DriverApi& d = driver();
  DCNN_REQUIRE(d.ok, "multicast driver API not available");
  cudaFree(0);
  CUmulticastObjectProp prop = mc_prop(world, (size_t)bytes);
  size_t g_mc = 0, g_mem = 0;
  DCNN_CHECK_CU(d.MulticastGetGranularity(&g_mc, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  CUmemAllocationProp ap;
  memset(&ap, 0, sizeof(ap));
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  DCNN_CHECK_CU(d.MemGetAllocationGranularity(&g_mem, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  const size_t g = g_mc > g_mem ? g_mc : g_mem;
  return (long long)(((size_t)bytes + g - 1) / g * g);
}

// Rank 0: create the multicast object for `world` devices and `bytes` (already padded); returns the generic handle
// (as an integer) and a POSIX file descriptor to pass to the other ranks.
DCNN_API int dcnn_nvls_create(int world, long long bytes, unsigned long long* handle_out, int* fd_out) {
  const DriverApi& d = driver();
  DCNN_REQUIRE(d.ok, "multicast driver API not available");
  cudaFree(0);
  CUmulticastObjectProp prop = mc_prop(world, (size_t)bytes);
  CUmemGenericAllocationHandle h;
  DCNN_CHECK_CU(d.MulticastCreate(&h, &prop));
  int fd = -1;
  DCNN_CHECK_CU(d.MemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *handle_out = (unsigned long long)h;
  *fd_out = fd;
  return 0;
}

// Other ranks: import the multicast object from the descriptor received from rank 0 (the descriptor is closed here).
DCNN_API int dcnn_nvls_import(int fd, unsigned long long* handle_out) {
  const DriverApi& d = driver();
  DCNN_REQUIRE(d.ok, "multicast driver API not available");
  cudaFree(0);
  CUmemGenericAllocationHandle h;
  DCNN_CHECK_CU(d.MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  close(fd);
  *handle_out = (unsigned long long)h;
  return 0;
}

// Every rank, before ANY rank binds memory (the caller puts a barrier between the two steps).
DCNN_API int dcnn_nvls_add_device(unsigned long long handle, int device) {
  const DriverApi& d = driver();
  CUdevice dev;
  DCNN_CHECK_CU(d.DeviceGet(&dev, device));
  DCNN_CHECK_CU(d.MulticastAddDevice((CUmemGenericAllocationHandle)handle, dev));
  return 0;
}

// Every rank: allocate `bytes` of device memory, map it (unicast pointer), bind it to the multicast object at offset 0
// and map the multicast object (multicast pointer).  The memory is zero-initialised.  *mem_handle_out is the allocation
// handle dcnn_nvls_release needs.  Whatever was set up before a failing step is undone.
DCNN_API int dcnn_nvls_bind(unsigned long long handle, int device, long long bytes, void** uc_ptr, void** mc_ptr,
                            unsigned long long* mem_handle_out) {
  const DriverApi& d = driver();
  DCNN_REQUIRE(d.ok, "multicast driver API not available");
  const size_t size = (size_t)bytes;
  CUmemAllocationProp ap;
  memset(&ap, 0, sizeof(ap));
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle mem = 0;
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CUdeviceptr uc = 0, mc = 0;
  bool uc_mapped = false, mc_mapped = false;
  CUresult r = d.MemCreate(&mem, size, &ap, 0);
  const char* what = "cuMemCreate";
  if (r == CUDA_SUCCESS) { what = "cuMemAddressReserve"; r = d.MemAddressReserve(&uc, size, 0, 0, 0); }
  if (r == CUDA_SUCCESS) { what = "cuMemMap"; r = d.MemMap(uc, size, 0, mem, 0); uc_mapped = r == CUDA_SUCCESS; }
  if (r == CUDA_SUCCESS) { what = "cuMemSetAccess"; r = d.MemSetAccess(uc, size, &acc, 1); }
  if (r == CUDA_SUCCESS) {
    what = "cudaMemset";
    if (cudaMemset((void*)uc, 0, size) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) r = CUDA_ERROR_UNKNOWN;
  }
  if (r == CUDA_SUCCESS) { what = "cuMulticastBindMem"; r = d.MulticastBindMem((CUmemGenericAllocationHandle)handle, 0, mem, 0, size, 0); }
  if (r == CUDA_SUCCESS) { what = "cuMemAddressReserve (multicast)"; r = d.MemAddressReserve(&mc, size, 0, 0, 0); }
  if (r == CUDA_SUCCESS) {
    what = "cuMemMap (multicast)";
    r = d.MemMap(mc, size, 0, (CUmemGenericAllocationHandle)handle, 0);
    mc_mapped = r == CUDA_SUCCESS;
  }
  if (r == CUDA_SUCCESS) { what = "cuMemSetAccess (multicast)"; r = d.MemSetAccess(mc, size, &acc, 1); }
  if (r != CUDA_SUCCESS) {
    if (mc_mapped) d.MemUnmap(mc, size);
    if (mc) d.MemAddressFree(mc, size);
    if (uc_mapped) d.MemUnmap(uc, size);
    if (uc) d.MemAddressFree(uc, size);
    if (mem) d.MemRelease(mem);
    cudaGetLastError();
    DCNN_REQUIRE(false, "dcnn_nvls_bind: %s failed with CUresult %d", what, (int)r);
  }
  *uc_ptr = (void*)uc;
  *mc_ptr = (void*)mc;
  *mem_handle_out = (unsigned long long)mem;
  return 0;
}

// Undo dcnn_nvls_bind (and drop this process's reference to the multicast object): unmap and free both address ranges,
// release the allocation and the object handle.  Every rank calls it once all work on the buffers has finished.
DCNN_API int dcnn_nvls_release(unsigned long long handle, unsigned long long mem_handle, void* uc_ptr, void* mc_ptr,
                               long long bytes) {
  const DriverApi& d = driver();
  DCNN_REQUIRE(d.ok, "multicast driver API not available");
  const size_t size = (size_t)bytes;
  DCNN_CHECK_CUDA(cudaDeviceSynchronize());
  if (mc_ptr) {
    DCNN_CHECK_CU(d.MemUnmap((CUdeviceptr)mc_ptr, size));
    DCNN_CHECK_CU(d.MemAddressFree((CUdeviceptr)mc_ptr, size));
  }
  if (uc_ptr) {
    DCNN_CHECK_CU(d.MemUnmap((CUdeviceptr)uc_ptr, size));
    DCNN_CHECK_CU(d.MemAddressFree((CUdeviceptr)uc_ptr, size));
  }
  if (mem_handle) DCNN_CHECK_CU(d.MemRelease((CUmemGenericAllocationHandle)mem_handle));
  if (handle) DCNN_CHECK_CU(d.MemRelease((CUmemGenericAllocationHandle)handle));
  return 0;
}

// mc_grads / mc_params: multicast pointers; params / grads: this replica's unicast pointers of the same buffers.
DCNN_API int dcnn_fused_allreduce_update_nvls(const float* mc_grads, float* mc_params, const float* params, float* grads,
                                              float* s1, float* s2, long long n, long long shard_begin,
                                              long long shard_end, unsigned long long* const* start_flag_ptrs,
                                              unsigned long long* const* done_flag_ptrs, unsigned int* block_counter,
                                              int rank, int world, unsigned long long* epoch_counter, int mode, float lr,
                                              float a, float b, float eps, float bc1, float bc2, float wd,
                                              const float* hyper, int num_blocks, cudaStream_t stream) {
  DCNN_REQUIRE((shard_begin & 3) == 0 && (shard_end & 3) == 0 && (n & 3) == 0,
               "fused all-reduce needs 4-element aligned shards");
  DCNN_REQUIRE(world >= 1 && world <= kMaxPeers, "world size %d out of range (max %d)", world, kMaxPeers);
  NvlsParams p;
  memset(&p, 0, sizeof(p));
  p.mc_grads = mc_grads; p.mc_params = mc_params; p.params = params; p.s1 = s1; p.s2 = s2;
  p.shard_begin = shard_begin; p.shard_end = shard_end;
  for (int r = 0; r < kMaxPeers; ++r) {
    p.start_flags.flags[r] = r < world ? start_flag_ptrs[r] : nullptr;
    p.done_flags.flags[r] = r < world ? done_flag_ptrs[r] : nullptr;
  }
  p.block_counter = block_counter; p.rank = rank; p.world = world; p.epoch_ptr = epoch_counter; p.hyper = hyper;
  p.lr = lr; p.a = a; p.b = b; p.eps = eps; p.bc1 = bc1; p.bc2 = bc2; p.wd = wd; p.inv_world = 1.f / (float)world;
  p.mode = mode;
  if (num_blocks <= 0) num_blocks = num_sms();
  epoch_bump_kernel<<<1, 1, 0, stream>>>(epoch_counter);
  fused_allreduce_update_nvls_kernel<<<num_blocks, 512, 0, stream>>>(p);
  DCNN_CHECK_LAUNCH();
  nvls_wait_done_zero_kernel<<<num_sms(), 512, 0, stream>>>(p.done_flags, rank, world, epoch_counter, grads, n);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(3);
  return 0;
}
