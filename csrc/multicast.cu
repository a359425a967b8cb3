// NVLS (NVSwitch multicast / in-switch reduction) variants of the model exchange.
//
// reference sites these stand in for: gossip_model_stage.py:85-132 (one finished model copied to every
// neighbour: here ONE multimem.st reaches all GPUs) and fedavg.py:49-58 for the special case "the
// train set is every GPU of a full mesh" (a weighted all-reduce: multimem.ld_reduce sums the
// contributions inside the switch).  SURVEY.md section 5.8 lists both as the multicast variants of G3
// and as the upper-bound comparison for config 5; the general product path stays the per-peer
// push + FedAvg kernel of gossip.cu, which handles subsets, tables and partial aggregates.
//
// A multicast object (cuMulticastCreate) spans N GPUs; each rank binds a physical allocation of
// its own to it.  A store to the multicast address lands in every rank's allocation; a
// multimem.ld_reduce of it returns the sum of all ranks' values.  The object is created by one
// process, exported as a POSIX file descriptor and imported by the others (the Python side passes
// the descriptor over a unix socket); in one process with several GPUs the handle is used directly.
// Availability is probed at run time (CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED); callers fall back
// to the unicast kernels when it is missing.
#include <cooperative_groups.h>
#include <cuda.h>
#include <unistd.h>

#include <mutex>

#include "common.cuh"

using namespace p2pfl;
namespace cg = cooperative_groups;

namespace {

template <typename F>
F driver_fn(const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
  return reinterpret_cast<F>(p);
}

#define DRV(type, name) static auto name##_fn = driver_fn<type>(#name)

typedef CUresult (*MulticastCreateFn)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
typedef CUresult (*MulticastAddDeviceFn)(CUmemGenericAllocationHandle, CUdevice);
typedef CUresult (*MulticastBindMemFn)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                                       unsigned long long);
typedef CUresult (*MulticastGranFn)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
typedef CUresult (*MemCreateFn)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
typedef CUresult (*MemReleaseFn)(CUmemGenericAllocationHandle);
typedef CUresult (*MemReserveFn)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
typedef CUresult (*MemMapFn)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
typedef CUresult (*MemSetAccessFn)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
typedef CUresult (*MemUnmapFn)(CUdeviceptr, size_t);
typedef CUresult (*MemFreeAddrFn)(CUdeviceptr, size_t);
typedef CUresult (*MemExportFn)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
typedef CUresult (*MemImportFn)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
typedef CUresult (*MemGranFn)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
typedef CUresult (*DeviceGetFn)(CUdevice*, int);
typedef CUresult (*DeviceGetAttrFn)(int*, CUdevice_attribute, CUdevice);

#define DRV_OK(expr)                                                                     \
  do {                                                                                   \
    CUresult _r = (expr);                                                                \
    if (_r != CUDA_SUCCESS) {                                                            \
      ::p2pfl::set_error("%s failed with CUresult %d (%s:%d)", #expr, (int)_r, __FILE__, __LINE__); \
      return -(int)_r - 1000;                                                            \
    }                                                                                    \
  } while (0)

struct Group {
  CUmemGenericAllocationHandle mc = 0;
  size_t size = 0;
  int n_devices = 0;
};
struct Binding {
  CUmemGenericAllocationHandle phys = 0;
  CUdeviceptr local = 0, mcva = 0;
  size_t size = 0;
};

CUmulticastObjectProp mc_prop(int n_devices, size_t size) {
  CUmulticastObjectProp p = {};
  p.numDevices = (unsigned)n_devices;
  p.size = size;
  p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  p.flags = 0;
  return p;
}

// ----------------------------------------------------------------- device code
__device__ __forceinline__ void mm_red_add_release(uint32_t* mc_flag, uint32_t v) {
  asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc_flag), "r"(v) : "memory");
}
__device__ __forceinline__ float4 mm_ld_reduce_f32x4(const float* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ void mm_st_f32x4(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

// every rank arrives on the multicast flag (one add lands in every rank's copy) and waits on its own copy
__device__ __forceinline__ int rank_barrier(cg::grid_group& grid, uint32_t* mc_flag, const uint32_t* local_flag,
                                            uint32_t target, uint64_t timeout_cycles) {
  __threadfence_system();
  grid.sync();
  __shared__ int s_ok;
  if (blockIdx.x == 0 && threadIdx.x == 0) mm_red_add_release(mc_flag, 1u);
  if (threadIdx.x == 0) {
    const long long start = clock64();
    int ok = 1;
    while ((int32_t)(ld_acquire_sys_u32(local_flag) - target) < 0) {
      if (timeout_cycles && (uint64_t)(clock64() - start) > timeout_cycles) {
        ok = 0;
        break;
      }
      __nanosleep(64);
    }
    s_ok = ok;
  }
  __syncthreads();
  return s_ok;
}

// Two-shot weighted all-reduce through the switch.  `local` is this rank's allocation bound to the
// multicast object, `mc` the multicast address of the same bytes; flags[0..1] live in the same
// allocation (`local_flags` / `mc_flags`).  On return params = sum_r w_r * params_r on every rank.
__global__ void __launch_bounds__(512)
mc_allreduce_kernel(float* __restrict__ params, __nv_bfloat16* __restrict__ params_bf16, float* __restrict__ local,
                    float* __restrict__ mc, uint32_t* mc_flags, const uint32_t* local_flags, uint64_t n, int rank, int n_ranks,
                    float weight, uint32_t epoch, uint64_t timeout_cycles, int32_t* status) {
  cg::grid_group grid = cg::this_grid();
  const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t n4 = n / 4;
  // 1. this rank's contribution, already weighted, into its bound allocation
  for (uint64_t i = tid; i < n4; i += nth) {
    float4 v = reinterpret_cast<const float4*>(params)[i];
    v.x *= weight;
    v.y *= weight;
    v.z *= weight;
    v.w *= weight;
    reinterpret_cast<float4*>(local)[i] = v;
  }
  int ok = rank_barrier(grid, mc_flags + 0, local_flags + 0, (uint32_t)n_ranks * epoch, timeout_cycles);
  // 2. this rank's slice: the switch returns the sum over all ranks, one store puts it everywhere
  const uint64_t per = (n4 + n_ranks - 1) / n_ranks;
  const uint64_t lo = per * rank, hi = (lo + per) < n4 ? (lo + per) : n4;
  if (ok) {
    for (uint64_t i = lo + tid; i < hi; i += nth) {
      const float4 s = mm_ld_reduce_f32x4(mc + 4 * i);
      mm_st_f32x4(mc + 4 * i, s);
    }
  }
  ok &= rank_barrier(grid, mc_flags + 1, local_flags + 1, (uint32_t)n_ranks * epoch, timeout_cycles);
  // 3. install
  if (ok) {
    for (uint64_t i = tid; i < n4; i += nth) {
      const uint4 u = ld_coherent_u4(local + 4 * i);  // written by the switch: not through L1
      const float4 v = make_float4(__uint_as_float(u.x), __uint_as_float(u.y), __uint_as_float(u.z), __uint_as_float(u.w));
      reinterpret_cast<float4*>(params)[i] = v;
      if (params_bf16) reinterpret_cast<uint2*>(params_bf16)[i] = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
    }
  } else if (tid == 0 && status) {
    atomicExch(status, 1);
  }
}

// G3 through the switch: one multimem.st per 16 bytes reaches every GPU's copy; then the flag.
__global__ void __launch_bounds__(512)
mc_store_kernel(const float* __restrict__ src, float* __restrict__ mc, uint32_t* mc_flag, uint64_t n, uint32_t seq) {
  cg::grid_group grid = cg::this_grid();
  const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = tid; i < n / 4; i += nth) mm_st_f32x4(mc + 4 * i, reinterpret_cast<const float4*>(src)[i]);
  __threadfence_system();
  grid.sync();
  if (tid == 0) asm volatile("multimem.st.release.sys.global.u32 [%0], %1;" ::"l"(mc_flag), "r"(seq) : "memory");
}

int coop_grid(int device, const void* fn, int threads) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, threads, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
  if (per_sm > 2) per_sm = 2;
  return per_sm * sm_count(device);
}

}  // namespace

extern "C" {

int p2pfl_mc_supported(int device) {
  DRV(DeviceGetFn, cuDeviceGet);
  DRV(DeviceGetAttrFn, cuDeviceGetAttribute);
  DRV(MulticastCreateFn, cuMulticastCreate);
  if (!cuDeviceGet_fn || !cuDeviceGetAttribute_fn || !cuMulticastCreate_fn) return 0;
  if (cudaSetDevice(device) != cudaSuccess) return 0;
  cudaFree(0);
  CUdevice d;
  int yes = 0;
  if (cuDeviceGet_fn(&d, device) != CUDA_SUCCESS) return 0;
  if (cuDeviceGetAttribute_fn(&yes, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d) != CUDA_SUCCESS) return 0;
  return yes ? 1 : 0;
}

/* size of a multicast object / bound allocation for `bytes` of payload over n_devices GPUs */
int p2pfl_mc_rounded_size(int n_devices, uint64_t bytes, uint64_t* out) {
  DRV(MulticastGranFn, cuMulticastGetGranularity);
  P2PFL_REQUIRE(cuMulticastGetGranularity_fn != nullptr && out != nullptr, "multicast is not available from this driver");
  CUmulticastObjectProp p = mc_prop(n_devices, bytes);
  size_t g = 0;
  DRV_OK(cuMulticastGetGranularity_fn(&g, &p, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  *out = (bytes + g - 1) / g * g;
  return 0;
}

/* creator side: the object, and a POSIX file descriptor other processes can import it from */
int p2pfl_mc_create(int device, int n_devices, uint64_t size, void** out_group, int* out_fd) {
  DRV(MulticastCreateFn, cuMulticastCreate);
  DRV(MemExportFn, cuMemExportToShareableHandle);
  P2PFL_REQUIRE(cuMulticastCreate_fn && cuMemExportToShareableHandle_fn && out_group, "multicast is not available from this driver");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  cudaFree(0);
  CUmulticastObjectProp p = mc_prop(n_devices, size);
  Group* g = new Group();
  g->size = size;
  g->n_devices = n_devices;
  CUresult r = cuMulticastCreate_fn(&g->mc, &p);
  if (r != CUDA_SUCCESS) {
    delete g;
    set_error("cuMulticastCreate failed with CUresult %d (%d devices, %llu bytes)", (int)r, n_devices, (unsigned long long)size);
    return -(int)r - 1000;
  }
  if (out_fd) {
    int fd = -1;
    r = cuMemExportToShareableHandle_fn(&fd, g->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
      set_error("cuMemExportToShareableHandle failed with CUresult %d", (int)r);
      return -(int)r - 1000;
    }
    *out_fd = fd;
  }
  *out_group = g;
  return 0;
}

int p2pfl_mc_import(int device, int fd, int n_devices, uint64_t size, void** out_group) {
  DRV(MemImportFn, cuMemImportFromShareableHandle);
  P2PFL_REQUIRE(cuMemImportFromShareableHandle_fn && out_group, "multicast is not available from this driver");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  cudaFree(0);
  Group* g = new Group();
  g->size = size;
  g->n_devices = n_devices;
  CUresult r = cuMemImportFromShareableHandle_fn(&g->mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) {
    delete g;
    set_error("cuMemImportFromShareableHandle failed with CUresult %d", (int)r);
    return -(int)r - 1000;
  }
  *out_group = g;
  return 0;
}

int p2pfl_mc_add_device(void* group, int device) {
  DRV(MulticastAddDeviceFn, cuMulticastAddDevice);
  DRV(DeviceGetFn, cuDeviceGet);
  P2PFL_REQUIRE(group && cuMulticastAddDevice_fn && cuDeviceGet_fn, "bad group");
  CUdevice d;
  DRV_OK(cuDeviceGet_fn(&d, device));
  DRV_OK(cuMulticastAddDevice_fn(static_cast<Group*>(group)->mc, d));
  return 0;
}

/* after EVERY device was added: allocate this device's memory, bind it, map both views */
int p2pfl_mc_bind(void* group, int device, void** out_binding, void** out_local, void** out_mc) {
  DRV(MemCreateFn, cuMemCreate);
  DRV(MemReserveFn, cuMemAddressReserve);
  DRV(MemMapFn, cuMemMap);
  DRV(MemSetAccessFn, cuMemSetAccess);
  DRV(MulticastBindMemFn, cuMulticastBindMem);
  DRV(MemGranFn, cuMemGetAllocationGranularity);
  P2PFL_REQUIRE(group && out_binding && out_local && out_mc, "null argument");
  P2PFL_REQUIRE(cuMemCreate_fn && cuMemAddressReserve_fn && cuMemMap_fn && cuMemSetAccess_fn && cuMulticastBindMem_fn &&
                    cuMemGetAllocationGranularity_fn,
                "virtual memory management is not available from this driver");
  Group* g = static_cast<Group*>(group);
  P2PFL_CUDA_OK(cudaSetDevice(device));
  cudaFree(0);
  CUmemAllocationProp prop = {};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  DRV_OK(cuMemGetAllocationGranularity_fn(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  P2PFL_REQUIRE(g->size % gran == 0, "multicast size is not a multiple of the allocation granularity");
  Binding* b = new Binding();
  b->size = g->size;
  DRV_OK(cuMemCreate_fn(&b->phys, b->size, &prop, 0));
  DRV_OK(cuMulticastBindMem_fn(g->mc, 0, b->phys, 0, b->size, 0));
  CUmemAccessDesc acc = {};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  DRV_OK(cuMemAddressReserve_fn(&b->local, b->size, gran, 0, 0));
  DRV_OK(cuMemMap_fn(b->local, b->size, 0, b->phys, 0));
  DRV_OK(cuMemSetAccess_fn(b->local, b->size, &acc, 1));
  DRV_OK(cuMemAddressReserve_fn(&b->mcva, b->size, gran, 0, 0));
  DRV_OK(cuMemMap_fn(b->mcva, b->size, 0, g->mc, 0));
  DRV_OK(cuMemSetAccess_fn(b->mcva, b->size, &acc, 1));
  P2PFL_CUDA_OK(cudaMemset(reinterpret_cast<void*>(b->local), 0, b->size));
  P2PFL_CUDA_OK(cudaDeviceSynchronize());
  *out_binding = b;
  *out_local = reinterpret_cast<void*>(b->local);
  *out_mc = reinterpret_cast<void*>(b->mcva);
  return 0;
}

int p2pfl_mc_unbind(void* binding) {
  DRV(MemUnmapFn, cuMemUnmap);
  DRV(MemFreeAddrFn, cuMemAddressFree);
  DRV(MemReleaseFn, cuMemRelease);
  Binding* b = static_cast<Binding*>(binding);
  if (!b) return 0;
  if (cuMemUnmap_fn && cuMemAddressFree_fn && cuMemRelease_fn) {
    cudaDeviceSynchronize();
    if (b->mcva) {
      cuMemUnmap_fn(b->mcva, b->size);
      cuMemAddressFree_fn(b->mcva, b->size);
    }
    if (b->local) {
      cuMemUnmap_fn(b->local, b->size);
      cuMemAddressFree_fn(b->local, b->size);
    }
    if (b->phys) cuMemRelease_fn(b->phys);
  }
  delete b;
  return 0;
}

int p2pfl_mc_destroy(void* group) {
  DRV(MemReleaseFn, cuMemRelease);
  Group* g = static_cast<Group*>(group);
  if (!g) return 0;
  if (cuMemRelease_fn && g->mc) cuMemRelease_fn(g->mc);
  delete g;
  return 0;
}

/* Weighted all-reduce of `params` (fp32[n], n % 4 == 0) over the ranks of the multicast group.
 * `local` / `mc`: the two views of this rank's bound allocation, at least n*4 + 256 bytes; the last
 * 256 bytes hold the two barrier counters.  `epoch` = 1, 2, 3, ... (counters are never reset). */
int p2pfl_mc_allreduce(int device, void* stream, float* params, void* params_bf16, void* local, void* mc, uint64_t n,
                       uint64_t flags_offset_bytes, int rank, int n_ranks, float weight, uint32_t epoch,
                       uint64_t timeout_cycles, int32_t* status) {
  P2PFL_REQUIRE(params && local && mc && n > 0 && n % 4 == 0 && n_ranks >= 1 && rank >= 0 && rank < n_ranks, "bad arguments");
  P2PFL_REQUIRE(flags_offset_bytes >= n * 4 && flags_offset_bytes % 16 == 0, "flags must follow the payload");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  float* lp = static_cast<float*>(local);
  float* mp = static_cast<float*>(mc);
  uint32_t* lf = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(local) + flags_offset_bytes);
  uint32_t* mf = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(mc) + flags_offset_bytes);
  __nv_bfloat16* pb = static_cast<__nv_bfloat16*>(params_bf16);
  const int grid = coop_grid(device, (const void*)mc_allreduce_kernel, 512);
  void* args[] = {&params, &pb, &lp, &mp, &mf, &lf, &n, &rank, &n_ranks, &weight, &epoch, &timeout_cycles, &status};
  P2PFL_CUDA_OK(cudaLaunchCooperativeKernel((void*)mc_allreduce_kernel, dim3(grid), dim3(512), args, 0, as_stream(stream)));
  return 0;
}

/* G3 through the switch: src (fp32[n]) lands in EVERY rank's bound allocation, then flag = seq. */
int p2pfl_mc_store(int device, void* stream, const float* src, void* mc, uint64_t n, uint64_t flag_offset_bytes, uint32_t seq) {
  P2PFL_REQUIRE(src && mc && n > 0 && n % 4 == 0 && flag_offset_bytes >= n * 4, "bad arguments");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  float* mp = static_cast<float*>(mc);
  uint32_t* mf = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(mc) + flag_offset_bytes);
  const int grid = coop_grid(device, (const void*)mc_store_kernel, 512);
  void* args[] = {&src, &mp, &mf, &n, &seq};
  P2PFL_CUDA_OK(cudaLaunchCooperativeKernel((void*)mc_store_kernel, dim3(grid), dim3(512), args, 0, as_stream(stream)));
  return 0;
}

}  // extern "C"
