// Gossip data plane, bulk-copy (TMA engine) variants of the model push.
//
// reference sites replaced: gossiper.py:239 (client.send of a model) and, for the remote-reduce
// mode, fedavg.py:49-54 (the weighted sum) - SURVEY.md section 5.8, kernel G1 modes A and B, and G3.
//
// gossip.cu moves tiles with 16-byte loads and stores issued by 512 threads per CTA.  Here the
// copy engine does the moving: a tile of the local model is fetched with ONE cp.async.bulk
// (global -> shared, completion on an mbarrier), converted in shared memory if the wire is bf16,
// and leaves with one bulk operation per peer:
//   mode B / G3  cp.async.bulk.global.shared::cta            store into the peer's inbox slot
//                                                            (or straight into its parameters)
//   mode A       cp.reduce.async.bulk.global.shared::cta.add.f32
//                                                            the peer's L2 adds w_i * tile into its
//                                                            accumulator: no receiver SM touches it
// The threads only convert / scale and publish the arrival flags (st.release.sys after the bulk
// group has completed); two tile buffers per CTA keep a load, a conversion and the stores in flight.
// All bulk LOADS read this GPU's memory only; the stores and reductions target peer-mapped memory.
#include "common.cuh"

using namespace p2pfl;

namespace {

constexpr int kThreads = 256;
constexpr uint32_t kTile = 16384;  // elements per bulk tile = per arrival flag: 64 KB fp32 in, 32 KB bf16 out
constexpr int kStages = 2;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
  }
}
// global (this GPU) -> shared, completion counted on the mbarrier
__device__ __forceinline__ void bulk_load(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// shared -> global (possibly a peer GPU's memory), tracked by the thread's bulk async-group
__device__ __forceinline__ void bulk_store(void* dst, uint32_t src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src_smem), "r"(bytes) : "memory");
}
// shared -> global with an fp32 add performed at the destination
__device__ __forceinline__ void bulk_reduce_add_f32(void* dst, uint32_t src_smem, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst), "r"(src_smem),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all committed groups have completed (their writes are performed), not merely read their source
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

struct BulkArgs {
  const float* src;      // local fp32 model
  uint64_t n;            // elements of the model
  uint64_t begin, end;   // element range to push (multiples of kTile, or end == n)
  uint32_t seq;
  int32_t mode;          // 0: store wire tiles into inbox slots (B / G3), 1: remote fp32 reduce (A)
  float scale;           // mode 1: FedAvg weight of this sender
};

// One CTA walks tiles begin/kTile + blockIdx.x, + gridDim.x, ...  Thread 0 drives the copy engine.
__global__ void __launch_bounds__(kThreads)
gossip_push_bulk_kernel(const BulkArgs a, const __grid_constant__ p2pfl_push_table t) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[kStages];
  __shared__ p2pfl_push_table s_t;
  float* in[kStages];
  uint8_t* out[kStages];
  for (int s = 0; s < kStages; ++s) {
    in[s] = reinterpret_cast<float*>(smem + (size_t)s * (kTile * 4 + kTile * 2));
    out[s] = reinterpret_cast<uint8_t*>(in[s]) + kTile * 4;
  }
  if (threadIdx.x == 0) {
    s_t = t;
    for (int s = 0; s < kStages; ++s) mbar_init(smem_u32(&full[s]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const bool bf16 = a.mode == 0 && s_t.wire_bf16 != 0;
  const uint64_t first = a.begin / kTile;
  const uint64_t n_tiles = (a.end - a.begin + kTile - 1) / kTile;
  const uint64_t mine = n_tiles > blockIdx.x ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  auto tile_len = [&](uint64_t k) {
    const uint64_t e0 = (first + blockIdx.x + k * gridDim.x) * kTile;
    return (uint32_t)((a.end - e0) < kTile ? (a.end - e0) : kTile);
  };
  auto issue_load = [&](uint64_t k) {
    const int s = (int)(k % kStages);
    const uint64_t e0 = (first + blockIdx.x + k * gridDim.x) * kTile;
    const uint32_t bytes = tile_len(k) * 4;
    mbar_expect_tx(smem_u32(&full[s]), bytes);
    bulk_load(smem_u32(in[s]), a.src + e0, bytes, smem_u32(&full[s]));
  };
  if (threadIdx.x == 0 && mine > 0) issue_load(0);

  for (uint64_t k = 0; k < mine; ++k) {
    const int s = (int)(k % kStages);
    const uint64_t e0 = (first + blockIdx.x + k * gridDim.x) * kTile;
    const uint32_t len = tile_len(k);
    // the buffers of stage (k+1) % kStages were last used by tile k-1, whose stores thread 0 has
    // already waited for (bulk_wait_all below): the next load may start now
    if (threadIdx.x == 0 && k + 1 < mine) issue_load(k + 1);
    mbar_wait(smem_u32(&full[s]), (uint32_t)(k / kStages) & 1u);
    uint32_t src_smem;
    uint32_t bytes;
    if (bf16) {
      // fp32 -> bf16 in shared memory, 8 elements per thread and pass
      for (uint32_t e = threadIdx.x * 8; e < len; e += kThreads * 8) {
        const float4 x = *reinterpret_cast<const float4*>(in[s] + e);
        const float4 y = *reinterpret_cast<const float4*>(in[s] + e + 4);
        uint4 w;
        w.x = pack_bf16x2(x.x, x.y);
        w.y = pack_bf16x2(x.z, x.w);
        w.z = pack_bf16x2(y.x, y.y);
        w.w = pack_bf16x2(y.z, y.w);
        *reinterpret_cast<uint4*>(out[s] + (size_t)e * 2) = w;
      }
      src_smem = smem_u32(out[s]);
      bytes = len * 2;
    } else if (a.mode == 1) {
      for (uint32_t e = threadIdx.x * 4; e < len; e += kThreads * 4) {
        float4 x = *reinterpret_cast<float4*>(in[s] + e);
        x.x *= a.scale;
        x.y *= a.scale;
        x.z *= a.scale;
        x.w *= a.scale;
        *reinterpret_cast<float4*>(in[s] + e) = x;
      }
      src_smem = smem_u32(in[s]);
      bytes = len * 4;
    } else {
      src_smem = smem_u32(in[s]);
      bytes = len * 4;
    }
    fence_async();    // the threads' writes to shared memory vs. the copy engine's reads
    __syncthreads();
    if (threadIdx.x == 0) {
      const size_t wb = bf16 ? 2 : 4;
      for (int p = 0; p < s_t.n_peers; ++p) {
        uint8_t* dst = reinterpret_cast<uint8_t*>(s_t.inbox[p]) + e0 * wb;
        if (a.mode == 1) bulk_reduce_add_f32(dst, src_smem, bytes);
        else bulk_store(dst, src_smem, bytes);
      }
      bulk_commit();
      bulk_wait_all();  // performed at the peers: the flags may follow
      fence_async();
      // one arrival per (peer, tile): mode 1 counts the senders, mode 0 publishes `seq`
      const uint64_t ft = e0 / kTile;
      for (int p = 0; p < s_t.n_peers; ++p) {
        if (a.mode == 1) {
          asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(s_t.flags[p] + ft), "r"(1u) : "memory");
        } else {
          st_release_sys_u32(s_t.flags[p] + ft, a.seq);
        }
      }
    }
    __syncthreads();  // stage s is free for the load of tile k + kStages (issued by thread 0 next round)
  }
}

// Receiver side of mode A: per tile, wait until `expected` bulk tiles were added into `acc` by the
// peers' copy engines (counters in `counts`), then params = acc + w_self * params (+ bf16 copy), and
// leave acc and the counter zeroed for the next exchange.
__global__ void __launch_bounds__(kThreads)
gossip_finalize_acc_kernel(float* __restrict__ params, __nv_bfloat16* __restrict__ params_bf16, float* __restrict__ acc,
                           uint32_t* __restrict__ counts, uint64_t n, uint32_t expected, float w_self,
                           uint64_t timeout_cycles, int32_t* status) {
  constexpr uint32_t flag_tile = kTile;
  const uint64_t n_tiles = (n + flag_tile - 1) / flag_tile;
  __shared__ int s_ok;
  for (uint64_t ti = blockIdx.x; ti < n_tiles; ti += gridDim.x) {
    const uint64_t t0 = ti * flag_tile;
    const uint32_t len = (uint32_t)((n - t0) < flag_tile ? (n - t0) : flag_tile);
    const uint32_t want = expected;
    if (threadIdx.x == 0) {
      const long long start = clock64();
      int ok = 1;
      while (ld_acquire_sys_u32(counts + ti) < want) {
        if (timeout_cycles && (uint64_t)(clock64() - start) > timeout_cycles) {
          ok = 0;
          break;
        }
        __nanosleep(32);
      }
      s_ok = ok;
      if (!ok && status) atomicExch(status, 1);
    }
    __syncthreads();
    for (uint32_t e = threadIdx.x * 4; e < len; e += kThreads * 4) {
      // the adds were performed in this GPU's L2: read there, not through L1
      const uint4 av = ld_coherent_u4(acc + t0 + e);
      float4 p = *reinterpret_cast<float4*>(params + t0 + e);
      p.x = fmaf(w_self, p.x, __uint_as_float(av.x));
      p.y = fmaf(w_self, p.y, __uint_as_float(av.y));
      p.z = fmaf(w_self, p.z, __uint_as_float(av.z));
      p.w = fmaf(w_self, p.w, __uint_as_float(av.w));
      *reinterpret_cast<float4*>(params + t0 + e) = p;
      *reinterpret_cast<float4*>(acc + t0 + e) = make_float4(0.f, 0.f, 0.f, 0.f);
      if (params_bf16)
        *reinterpret_cast<uint2*>(params_bf16 + t0 + e) = make_uint2(pack_bf16x2(p.x, p.y), pack_bf16x2(p.z, p.w));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      counts[ti] = 0;
      if (status) atomicAdd(status + 1, 1);
    }
  }
}

int launch_bulk(int device, cudaStream_t stream, const BulkArgs& a, const p2pfl_push_table* table) {
  P2PFL_REQUIRE(table && table->n_peers >= 0 && table->n_peers <= P2PFL_MAX_PEERS, "bad push table");
  P2PFL_REQUIRE(a.src && a.n > 0 && a.begin < a.end && a.end <= a.n, "bad element range");
  P2PFL_REQUIRE(a.begin % kTile == 0 && (a.end % kTile == 0 || a.end == a.n), "ranges are whole tiles of 16384 elements");
  P2PFL_REQUIRE(a.n % 8 == 0, "model length must be a multiple of 8 elements");
  if (table->n_peers == 0) return 0;
  P2PFL_CUDA_OK(cudaSetDevice(device));
  constexpr int smem = kStages * (kTile * 4 + kTile * 2);
  static bool set[64] = {false};
  if (device >= 0 && device < 64 && !set[device]) {
    P2PFL_CUDA_OK(cudaFuncSetAttribute(gossip_push_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    set[device] = true;
  }
  const uint64_t tiles = (a.end - a.begin + kTile - 1) / kTile;
  const uint64_t cap = (uint64_t)sm_count(device) * 2;
  const int grid = (int)(tiles < cap ? tiles : cap);
  gossip_push_bulk_kernel<<<grid, kThreads, smem, stream>>>(a, *table);
  P2PFL_LAUNCH_OK();
  return 0;
}

}  // namespace

extern "C" {

int p2pfl_gossip_push_bulk(int device, void* stream, const float* src, uint64_t n, uint64_t begin, uint64_t end,
                           uint32_t seq, const p2pfl_push_table* table) {
  BulkArgs a;
  a.src = src;
  a.n = n;
  a.begin = begin;
  a.end = end;
  a.seq = seq;
  a.mode = 0;
  a.scale = 1.f;
  return launch_bulk(device, as_stream(stream), a, table);
}

int p2pfl_gossip_push_reduce_bulk(int device, void* stream, const float* src, uint64_t n, float weight,
                                  const p2pfl_push_table* table) {
  P2PFL_REQUIRE(table && table->wire_bf16 == 0, "the remote-reduce mode adds fp32 tiles");
  BulkArgs a;
  a.src = src;
  a.n = n;
  a.begin = 0;
  a.end = n;
  a.seq = 0;
  a.mode = 1;
  a.scale = weight;
  return launch_bulk(device, as_stream(stream), a, table);
}

int p2pfl_gossip_finalize_acc(int device, void* stream, float* params, void* params_bf16, float* acc, uint32_t* counts,
                              uint64_t n, int n_senders, float self_weight, uint64_t timeout_cycles, int32_t* status) {
  P2PFL_REQUIRE(params && acc && counts && n > 0 && n % 8 == 0 && n_senders >= 0, "bad arguments");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  const uint64_t tiles = (n + kTile - 1) / kTile;
  const uint64_t cap = (uint64_t)sm_count(device) * 4;
  const int grid = (int)(tiles < cap ? tiles : cap);
  gossip_finalize_acc_kernel<<<grid, kThreads, 0, as_stream(stream)>>>(
      params, reinterpret_cast<__nv_bfloat16*>(params_bf16), acc, counts, n, (uint32_t)n_senders, self_weight, timeout_cycles,
      status);
  P2PFL_LAUNCH_OK();
  return 0;
}

}  // extern "C"
