// Gossip data plane: model push over NVLink fused with FedAvg on arrival.
//
// reference sites replaced: gossiper.py:239 (client.send of a whole model),
// fedavg.py:49-58 (weighted sum on the CPU), lightning_learner.py:152
// (load_state_dict of the result).  See SURVEY.md Appendix C for the contract.
//
// Sender side   : read the local fp32 tile once, cast to the wire dtype, store
//                 it into slot[me] of EVERY peer (peer-mapped pointers; the
//                 stores travel over NVLink/NVSwitch), then publish `seq` in
//                 the tile's flag with st.release.sys.
// Receiver side : per tile, ld.acquire.sys-spin on the flags of the expected
//                 contributors, then accumulate w_i * tile_i in fp32 in slot
//                 order (deterministic), write the live parameters (+ bf16
//                 working copy).  A host-writable abort word and a cycle budget
//                 bound every wait; on abort a tile is finalised from the
//                 contributors that did arrive, renormalised.
//
// Senders never read remote memory; receivers never write remote memory.
#include <cooperative_groups.h>

#include "common.cuh"

using namespace p2pfl;

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ uint4 ld_volatile_u4(const void* p) {
  uint4 r;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

// ---------------------------------------------------------------------- push
// `cta` / `n_cta` let several logical ranks share one launch (emulation).
__device__ void push_phase(const float* __restrict__ src, uint64_t n, uint32_t tile, uint32_t seq,
                           const p2pfl_push_table& t, int cta, int n_cta) {
  const uint32_t n_tiles = (uint32_t)((n + tile - 1) / tile);
  for (uint32_t ti = cta; ti < n_tiles; ti += n_cta) {
    const uint64_t t0 = (uint64_t)ti * tile;
    const uint32_t len = (uint32_t)((n - t0) < tile ? (n - t0) : tile);
    if (t.wire_bf16) {
      // 8 elements per thread-iteration: 32 B in, 16 B out per peer
      for (uint32_t e = threadIdx.x * 8; e < len; e += kThreads * 8) {
        const float4 a = *reinterpret_cast<const float4*>(src + t0 + e);
        const float4 b = *reinterpret_cast<const float4*>(src + t0 + e + 4);
        uint4 w;
        w.x = pack_bf16x2(a.x, a.y);
        w.y = pack_bf16x2(a.z, a.w);
        w.z = pack_bf16x2(b.x, b.y);
        w.w = pack_bf16x2(b.z, b.w);
#pragma unroll 4
        for (int p = 0; p < t.n_peers; ++p)
          st_stream_u4(reinterpret_cast<__nv_bfloat16*>(t.inbox[p]) + t0 + e, w);
      }
    } else {
      for (uint32_t e = threadIdx.x * 4; e < len; e += kThreads * 4) {
        const uint4 w = *reinterpret_cast<const uint4*>(src + t0 + e);
#pragma unroll 4
        for (int p = 0; p < t.n_peers; ++p) st_stream_u4(reinterpret_cast<float*>(t.inbox[p]) + t0 + e, w);
      }
    }
    __syncthreads();  // every store of this tile happens-before the flag writers below
    if ((int)threadIdx.x < t.n_peers) {
      __threadfence_system();
      st_release_sys_u32(t.flags[threadIdx.x] + ti, seq);
    }
  }
}

// -------------------------------------------------------------------- reduce
__device__ void reduce_phase(float* __restrict__ params, __nv_bfloat16* __restrict__ params_bf16, uint64_t n,
                             uint32_t tile, uint32_t seq, const p2pfl_reduce_table& t,
                             const uint32_t* __restrict__ abort_flag, uint64_t timeout_cycles, int32_t* status, int cta,
                             int n_cta) {
  __shared__ int s_arrived[P2PFL_MAX_PEERS];
  const uint32_t n_tiles = (uint32_t)((n + tile - 1) / tile);
  for (uint32_t ti = cta; ti < n_tiles; ti += n_cta) {
    const uint64_t t0 = (uint64_t)ti * tile;
    const uint32_t len = (uint32_t)((n - t0) < tile ? (n - t0) : tile);

    // ---- wait: one thread per expected contributor ---------------------------
    if ((int)threadIdx.x < t.n_slots) {
      const uint32_t* f = t.flags[threadIdx.x] + ti;
      const long long start = clock64();
      int ok = 1;
      uint32_t spins = 0;
      while ((int32_t)(ld_acquire_sys_u32(f) - seq) < 0) {
        if ((++spins & 63u) == 0) {
          if ((abort_flag && ld_volatile_u32(abort_flag) != 0) ||
              (timeout_cycles && (uint64_t)(clock64() - start) > timeout_cycles)) {
            ok = 0;
            break;
          }
        }
        __nanosleep(64);
      }
      s_arrived[threadIdx.x] = ok;
      if (!ok && status) atomicExch(status, 1);
    }
    __syncthreads();

    // ---- weights of what actually arrived -----------------------------------
    float w[P2PFL_MAX_PEERS];
    float wsum = (t.self_pos >= 0) ? t.self_weight : 0.f;
#pragma unroll
    for (int j = 0; j < P2PFL_MAX_PEERS; ++j) {
      w[j] = (j < t.n_slots && s_arrived[j]) ? t.weight[j] : 0.f;
      wsum += w[j];
    }
    const float scale = (wsum > 0.f && fabsf(wsum - 1.f) > 1e-6f) ? 1.f / wsum : 1.f;
    const float w_self = (t.self_pos >= 0) ? t.self_weight : 0.f;

    // ---- accumulate in slot order, local model inserted at self_pos ---------
    for (uint32_t e = threadIdx.x * 8; e < len; e += kThreads * 8) {
      float own[8];
      {
        const float4 a = *reinterpret_cast<const float4*>(params + t0 + e);
        const float4 b = *reinterpret_cast<const float4*>(params + t0 + e + 4);
        own[0] = a.x; own[1] = a.y; own[2] = a.z; own[3] = a.w;
        own[4] = b.x; own[5] = b.y; own[6] = b.z; own[7] = b.w;
        if (t.round_self && t.wire_bf16) {
#pragma unroll
          for (int k = 0; k < 8; ++k) own[k] = __bfloat162float(__float2bfloat16_rn(own[k]));
        }
      }
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int j = 0; j <= t.n_slots; ++j) {
        if (j == t.self_pos) {
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] = fmaf(w_self, own[k], acc[k]);
        }
        if (j == t.n_slots) break;
        const float wj = w[j];
        if (wj == 0.f) continue;
        if (t.wire_bf16) {
          const uint4 u = ld_volatile_u4(reinterpret_cast<const __nv_bfloat16*>(t.slot[j]) + t0 + e);
          const float2 p0 = unpack_bf16x2(u.x), p1 = unpack_bf16x2(u.y), p2 = unpack_bf16x2(u.z), p3 = unpack_bf16x2(u.w);
          acc[0] = fmaf(wj, p0.x, acc[0]); acc[1] = fmaf(wj, p0.y, acc[1]);
          acc[2] = fmaf(wj, p1.x, acc[2]); acc[3] = fmaf(wj, p1.y, acc[3]);
          acc[4] = fmaf(wj, p2.x, acc[4]); acc[5] = fmaf(wj, p2.y, acc[5]);
          acc[6] = fmaf(wj, p3.x, acc[6]); acc[7] = fmaf(wj, p3.y, acc[7]);
        } else {
          const uint4 u0 = ld_volatile_u4(reinterpret_cast<const float*>(t.slot[j]) + t0 + e);
          const uint4 u1 = ld_volatile_u4(reinterpret_cast<const float*>(t.slot[j]) + t0 + e + 4);
          acc[0] = fmaf(wj, __uint_as_float(u0.x), acc[0]); acc[1] = fmaf(wj, __uint_as_float(u0.y), acc[1]);
          acc[2] = fmaf(wj, __uint_as_float(u0.z), acc[2]); acc[3] = fmaf(wj, __uint_as_float(u0.w), acc[3]);
          acc[4] = fmaf(wj, __uint_as_float(u1.x), acc[4]); acc[5] = fmaf(wj, __uint_as_float(u1.y), acc[5]);
          acc[6] = fmaf(wj, __uint_as_float(u1.z), acc[6]); acc[7] = fmaf(wj, __uint_as_float(u1.w), acc[7]);
        }
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] *= scale;
      *reinterpret_cast<float4*>(params + t0 + e) = make_float4(acc[0], acc[1], acc[2], acc[3]);
      *reinterpret_cast<float4*>(params + t0 + e + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
      if (params_bf16) {
        uint4 o;
        o.x = pack_bf16x2(acc[0], acc[1]);
        o.y = pack_bf16x2(acc[2], acc[3]);
        o.z = pack_bf16x2(acc[4], acc[5]);
        o.w = pack_bf16x2(acc[6], acc[7]);
        *reinterpret_cast<uint4*>(params_bf16 + t0 + e) = o;
      }
    }
    if (threadIdx.x == 0 && status) atomicAdd(status + 1, 1);
    __syncthreads();  // s_arrived is reused by the next tile
  }
}

// ------------------------------------------------------------------- kernels
__global__ void __launch_bounds__(kThreads)
gossip_push_kernel(const float* __restrict__ src, uint64_t n, uint32_t tile, uint32_t seq, p2pfl_push_table t) {
  push_phase(src, n, tile, seq, t, blockIdx.x, gridDim.x);
}

__global__ void __launch_bounds__(kThreads)
gossip_reduce_kernel(float* __restrict__ params, __nv_bfloat16* __restrict__ params_bf16, uint64_t n, uint32_t tile,
                     uint32_t seq, p2pfl_reduce_table t, const uint32_t* __restrict__ abort_flag,
                     uint64_t timeout_cycles, int32_t* status) {
  reduce_phase(params, params_bf16, n, tile, seq, t, abort_flag, timeout_cycles, status, blockIdx.x, gridDim.x);
}

// G1+G2: one launch.  Every CTA first pushes its tiles (no waiting), then
// reduces the same tile indices; peers run the same schedule, so tile i of
// every contributor is in flight while tile i-1 is being accumulated.
__global__ void __launch_bounds__(kThreads)
gossip_fused_kernel(float* __restrict__ params, __nv_bfloat16* __restrict__ params_bf16, uint64_t n, uint32_t tile,
                    uint32_t seq, p2pfl_push_table push, p2pfl_reduce_table red,
                    const uint32_t* __restrict__ abort_flag, uint64_t timeout_cycles, int32_t* status) {
  push_phase(params, n, tile, seq, push, blockIdx.x, gridDim.x);
  reduce_phase(params, params_bf16, n, tile, seq, red, abort_flag, timeout_cycles, status, blockIdx.x, gridDim.x);
}

// Test harness: R logical ranks in ONE cooperative launch on one GPU
// (blockIdx.y = rank).  Co-residency of all CTAs is what makes the cross-rank
// flag waits safe on a single device.
struct EmuArgs {
  float* params[P2PFL_MAX_PEERS];
  const p2pfl_push_table* push;     // device array [R]
  const p2pfl_reduce_table* reduce; // device array [R]
};

__global__ void __launch_bounds__(kThreads)
gossip_fused_emulated_kernel(EmuArgs a, uint64_t n, uint32_t tile, uint32_t seq, uint64_t timeout_cycles,
                             int32_t* status) {
  const int r = blockIdx.y;
  __shared__ p2pfl_push_table s_push;
  __shared__ p2pfl_reduce_table s_red;
  if (threadIdx.x == 0) {
    s_push = a.push[r];
    s_red = a.reduce[r];
  }
  __syncthreads();
  push_phase(a.params[r], n, tile, seq, s_push, blockIdx.x, gridDim.x);
  reduce_phase(a.params[r], nullptr, n, tile, seq, s_red, nullptr, timeout_cycles, status + 4 * r, blockIdx.x,
               gridDim.x);
}

int check_common(uint64_t n, uint32_t tile) {
  P2PFL_REQUIRE(n > 0 && (n % 8) == 0, "gossip kernels need n to be a positive multiple of 8 elements");
  P2PFL_REQUIRE(tile >= 2048 && (tile % 2048) == 0, "tile must be a multiple of 2048 elements");
  return 0;
}

int persistent_grid(int device, uint64_t n, uint32_t tile) {
  uint64_t tiles = (n + tile - 1) / tile;
  uint64_t cap = (uint64_t)sm_count(device) * 2;
  return (int)(tiles < cap ? tiles : cap);
}

}  // namespace

extern "C" {

int p2pfl_gossip_push(int device, void* stream, const float* src, uint64_t n, uint32_t tile, uint32_t seq,
                      const p2pfl_push_table* table) {
  if (int rc = check_common(n, tile)) return rc;
  P2PFL_REQUIRE(table && table->n_peers >= 0 && table->n_peers <= P2PFL_MAX_PEERS, "bad push table");
  if (table->n_peers == 0) return 0;
  P2PFL_CUDA_OK(cudaSetDevice(device));
  gossip_push_kernel<<<persistent_grid(device, n, tile), kThreads, 0, as_stream(stream)>>>(src, n, tile, seq, *table);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_gossip_reduce(int device, void* stream, float* params, void* params_bf16, uint64_t n, uint32_t tile,
                        uint32_t seq, const p2pfl_reduce_table* table, const uint32_t* abort_flag,
                        uint64_t timeout_cycles, int32_t* status) {
  if (int rc = check_common(n, tile)) return rc;
  P2PFL_REQUIRE(table && table->n_slots >= 0 && table->n_slots <= P2PFL_MAX_PEERS, "bad reduce table");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  gossip_reduce_kernel<<<persistent_grid(device, n, tile), kThreads, 0, as_stream(stream)>>>(
      params, (__nv_bfloat16*)params_bf16, n, tile, seq, *table, abort_flag, timeout_cycles, status);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_gossip_fused(int device, void* stream, float* params, void* params_bf16, uint64_t n, uint32_t tile,
                       uint32_t seq, const p2pfl_push_table* push, const p2pfl_reduce_table* reduce,
                       const uint32_t* abort_flag, uint64_t timeout_cycles, int32_t* status) {
  if (int rc = check_common(n, tile)) return rc;
  P2PFL_REQUIRE(push && reduce, "null table");
  P2PFL_REQUIRE(push->n_peers <= P2PFL_MAX_PEERS && reduce->n_slots <= P2PFL_MAX_PEERS, "too many peers");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  // one CTA per SM: the kernel spins on remote flags, keep it to a single wave
  uint64_t tiles = (n + tile - 1) / tile;
  int grid = (int)(tiles < (uint64_t)sm_count(device) ? tiles : (uint64_t)sm_count(device));
  gossip_fused_kernel<<<grid, kThreads, 0, as_stream(stream)>>>(params, (__nv_bfloat16*)params_bf16, n, tile, seq,
                                                               *push, *reduce, abort_flag, timeout_cycles, status);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_gossip_fused_emulated(int device, void* stream, int n_ranks, float* const* params, uint64_t n, uint32_t tile,
                                uint32_t seq, const p2pfl_push_table* push_tables,
                                const p2pfl_reduce_table* reduce_tables, uint64_t timeout_cycles, int32_t* status) {
  if (int rc = check_common(n, tile)) return rc;
  P2PFL_REQUIRE(n_ranks >= 1 && n_ranks <= P2PFL_MAX_PEERS, "bad rank count");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  EmuArgs a;
  for (int r = 0; r < n_ranks; ++r) a.params[r] = params[r];
  a.push = push_tables;
  a.reduce = reduce_tables;
  int per_sm = 0;
  P2PFL_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gossip_fused_emulated_kernel, kThreads, 0));
  int capacity = per_sm * sm_count(device);
  int gx = capacity / n_ranks;
  uint64_t tiles = (n + tile - 1) / tile;
  if ((uint64_t)gx > tiles) gx = (int)tiles;
  if (gx > 64) gx = 64;
  P2PFL_REQUIRE(gx >= 1, "not enough resident CTAs for the emulated exchange");
  dim3 grid(gx, n_ranks, 1), block(kThreads, 1, 1);
  void* args[] = {&a, &n, &tile, &seq, &timeout_cycles, &status};
  P2PFL_CUDA_OK(cudaLaunchCooperativeKernel((void*)gossip_fused_emulated_kernel, grid, block, args, 0, as_stream(stream)));
  return 0;
}

}  // extern "C"
