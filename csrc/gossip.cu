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

#include <mutex>

#include "common.cuh"

using namespace p2pfl;

namespace {

constexpr int kThreads = 512;
constexpr int kLag = 2;  // tiles between pushing a tile and reducing it (fused kernel)

// Inbox tiles are read only after the tile's flag was acquired at system scope and
// were never touched earlier in this launch (L1 is invalidated at launch), so a
// plain L1-bypassing load sees the peer's data; unlike ld.volatile it can be batched.
__device__ __forceinline__ uint4 ld_inbox_u4(const void* p) {
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

// ---------------------------------------------------------------------- push
// `snap` (nullable): the tile is also copied there in fp32, so that the host can put the
// pre-exchange model back if the exchange does not complete (the reduce overwrites `src`).
__device__ __forceinline__ void push_tile_data(const float* __restrict__ src, uint64_t n, uint32_t tile,
                                               const p2pfl_push_table& t, uint32_t ti, float* __restrict__ snap = nullptr) {
  const uint64_t t0 = (uint64_t)ti * tile;
  const uint32_t len = (uint32_t)((n - t0) < tile ? (n - t0) : tile);
  if (t.wire_bf16) {
    // 8 elements per vector (32 B in, 16 B out per peer), 4 vectors in flight per thread
    for (uint32_t e0 = threadIdx.x * 8; e0 < len; e0 += kThreads * 8 * 4) {
      uint4 w[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t e = e0 + u * kThreads * 8;
        if (e < len) {
          const float4 a = *reinterpret_cast<const float4*>(src + t0 + e);
          const float4 b = *reinterpret_cast<const float4*>(src + t0 + e + 4);
          if (snap) {
            *reinterpret_cast<float4*>(snap + t0 + e) = a;
            *reinterpret_cast<float4*>(snap + t0 + e + 4) = b;
          }
          w[u].x = pack_bf16x2(a.x, a.y);
          w[u].y = pack_bf16x2(a.z, a.w);
          w[u].z = pack_bf16x2(b.x, b.y);
          w[u].w = pack_bf16x2(b.z, b.w);
        }
      }
      for (int p = 0; p < t.n_peers; ++p) {
        __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(t.inbox[p]) + t0;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t e = e0 + u * kThreads * 8;
          if (e < len) st_stream_u4(dst + e, w[u]);
        }
      }
    }
  } else {
    for (uint32_t e0 = threadIdx.x * 4; e0 < len; e0 += kThreads * 4 * 4) {
      uint4 w[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t e = e0 + u * kThreads * 4;
        if (e < len) {
          w[u] = *reinterpret_cast<const uint4*>(src + t0 + e);
          if (snap) *reinterpret_cast<uint4*>(snap + t0 + e) = w[u];
        }
      }
      for (int p = 0; p < t.n_peers; ++p) {
        float* dst = reinterpret_cast<float*>(t.inbox[p]) + t0;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t e = e0 + u * kThreads * 4;
          if (e < len) st_stream_u4(dst + e, w[u]);
        }
      }
    }
  }
  __syncthreads();  // every store of this tile happens-before the flag writers
}

// Publish tile `ti`: st.release.sys orders every store this CTA made before the
// preceding barrier ahead of the flag (release is cumulative over bar.sync), so no
// separate __threadfence_system() is needed.
__device__ __forceinline__ void publish_tile(uint32_t seq, const p2pfl_push_table& t, uint32_t ti) {
  if ((int)threadIdx.x < t.n_peers) st_release_sys_u32(t.flags[threadIdx.x] + ti, seq);
}

// `cta` / `n_cta` let several logical ranks share one launch (emulation).
__device__ void push_phase(const float* __restrict__ src, uint64_t n, uint32_t tile, uint32_t seq,
                           const p2pfl_push_table& t, int cta, int n_cta) {
  const uint32_t n_tiles = (uint32_t)((n + tile - 1) / tile);
  for (uint32_t ti = cta; ti < n_tiles; ti += n_cta) {
    push_tile_data(src, n, tile, t, ti);
    publish_tile(seq, t, ti);
  }
}

// -------------------------------------------------------------------- reduce
__device__ __forceinline__ void reduce_tile(float* __restrict__ params, __nv_bfloat16* __restrict__ params_bf16,
                                            uint64_t n, uint32_t tile, uint32_t seq, const p2pfl_reduce_table& t,
                                            const uint32_t* __restrict__ abort_flag, uint64_t timeout_cycles,
                                            int32_t* status, uint32_t ti, int* s_arrived) {
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
      __nanosleep(32);
    }
    s_arrived[threadIdx.x] = ok;
    if (!ok && status) {
      atomicExch(status, 1);
      atomicOr(status + 2, 1 << threadIdx.x);  // which contributors (table positions) were missing somewhere
    }
  }
  __syncthreads();

  // ---- weights of what actually arrived -----------------------------------
  float wsum = (t.self_pos >= 0) ? t.self_weight : 0.f;
  for (int j = 0; j < t.n_slots; ++j) wsum += s_arrived[j] ? t.weight[j] : 0.f;
  const float scale = (wsum > 0.f && fabsf(wsum - 1.f) > 1e-6f) ? 1.f / wsum : 1.f;
  const float w_self = (t.self_pos >= 0) ? t.self_weight : 0.f;
  const bool round_own = t.round_self && t.wire_bf16;

  // ---- accumulate in slot order, local model inserted at self_pos ---------
  for (uint32_t e = threadIdx.x * 8; e < len; e += kThreads * 8) {
    float own[8], acc[8];
    {
      const float4 a = *reinterpret_cast<const float4*>(params + t0 + e);
      const float4 b = *reinterpret_cast<const float4*>(params + t0 + e + 4);
      own[0] = a.x; own[1] = a.y; own[2] = a.z; own[3] = a.w;
      own[4] = b.x; own[5] = b.y; own[6] = b.z; own[7] = b.w;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (round_own) own[k] = __bfloat162float(__float2bfloat16_rn(own[k]));
      acc[k] = 0.f;
    }
    for (int j = 0; j <= t.n_slots; ++j) {
      if (j == t.self_pos) {
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(w_self, own[k], acc[k]);
      }
      if (j == t.n_slots) break;
      const float wj = s_arrived[j] ? t.weight[j] : 0.f;
      if (wj == 0.f) continue;
      if (t.wire_bf16) {
        const uint4 v = ld_inbox_u4(reinterpret_cast<const __nv_bfloat16*>(t.slot[j]) + t0 + e);
        const float2 p0 = unpack_bf16x2(v.x), p1 = unpack_bf16x2(v.y), p2 = unpack_bf16x2(v.z), p3 = unpack_bf16x2(v.w);
        acc[0] = fmaf(wj, p0.x, acc[0]); acc[1] = fmaf(wj, p0.y, acc[1]);
        acc[2] = fmaf(wj, p1.x, acc[2]); acc[3] = fmaf(wj, p1.y, acc[3]);
        acc[4] = fmaf(wj, p2.x, acc[4]); acc[5] = fmaf(wj, p2.y, acc[5]);
        acc[6] = fmaf(wj, p3.x, acc[6]); acc[7] = fmaf(wj, p3.y, acc[7]);
      } else {
        const float* sp = reinterpret_cast<const float*>(t.slot[j]) + t0 + e;
        const uint4 v0 = ld_inbox_u4(sp), v1 = ld_inbox_u4(sp + 4);
        acc[0] = fmaf(wj, __uint_as_float(v0.x), acc[0]); acc[1] = fmaf(wj, __uint_as_float(v0.y), acc[1]);
        acc[2] = fmaf(wj, __uint_as_float(v0.z), acc[2]); acc[3] = fmaf(wj, __uint_as_float(v0.w), acc[3]);
        acc[4] = fmaf(wj, __uint_as_float(v1.x), acc[4]); acc[5] = fmaf(wj, __uint_as_float(v1.y), acc[5]);
        acc[6] = fmaf(wj, __uint_as_float(v1.z), acc[6]); acc[7] = fmaf(wj, __uint_as_float(v1.w), acc[7]);
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

__device__ void reduce_phase(float* __restrict__ params, __nv_bfloat16* __restrict__ params_bf16, uint64_t n,
                             uint32_t tile, uint32_t seq, const p2pfl_reduce_table& t,
                             const uint32_t* __restrict__ abort_flag, uint64_t timeout_cycles, int32_t* status, int cta,
                             int n_cta) {
  __shared__ int s_arrived[P2PFL_MAX_PEERS];
  const uint32_t n_tiles = (uint32_t)((n + tile - 1) / tile);
  for (uint32_t ti = cta; ti < n_tiles; ti += n_cta)
    reduce_tile(params, params_bf16, n, tile, seq, t, abort_flag, timeout_cycles, status, ti, s_arrived);
}

// Fused schedule of one CTA: push tile k, then reduce the tile pushed kLag steps
// earlier - peers run the same schedule, so by then their copy of that tile has
// normally landed and NVLink stores overlap the HBM-bound accumulate.
__device__ void fused_phase(float* __restrict__ params, __nv_bfloat16* __restrict__ params_bf16, uint64_t n,
                            uint32_t tile, uint32_t seq, const p2pfl_push_table& push, const p2pfl_reduce_table& red,
                            const uint32_t* __restrict__ abort_flag, uint64_t timeout_cycles, int32_t* status, int cta,
                            int n_cta, float* __restrict__ snap = nullptr) {
  __shared__ int s_arrived[P2PFL_MAX_PEERS];
  const uint32_t n_tiles = (uint32_t)((n + tile - 1) / tile);
  const uint32_t mine = n_tiles > (uint32_t)cta ? (n_tiles - cta + n_cta - 1) / n_cta : 0;
  for (uint32_t k = 0; k < mine + kLag; ++k) {
    if (k < mine) push_tile_data(params, n, tile, push, cta + k * n_cta, snap);
    if (k >= kLag)
      reduce_tile(params, params_bf16, n, tile, seq, red, abort_flag, timeout_cycles, status, cta + (k - kLag) * n_cta,
                  s_arrived);
    // flags go out AFTER the reduce work of this step: the tile's remote stores have had
    // that long to drain, so the release fence does not stall on an NVLink round trip
    if (k < mine) publish_tile(seq, push, cta + k * n_cta);
  }
}

// ------------------------------------------------------------------- kernels
__global__ void __launch_bounds__(kThreads)
gossip_push_kernel(const float* __restrict__ src, uint64_t n, uint32_t tile, uint32_t seq,
                   const __grid_constant__ p2pfl_push_table t) {
  // dynamic indexing into kernel parameters would be copied to local memory: stage in smem
  __shared__ p2pfl_push_table s_push;
  if (threadIdx.x == 0) s_push = t;
  __syncthreads();
  push_phase(src, n, tile, seq, s_push, blockIdx.x, gridDim.x);
}

__global__ void __launch_bounds__(kThreads, 2)
gossip_reduce_kernel(float* __restrict__ params, __nv_bfloat16* __restrict__ params_bf16, uint64_t n, uint32_t tile,
                     uint32_t seq, const __grid_constant__ p2pfl_reduce_table t,
                     const uint32_t* __restrict__ abort_flag, uint64_t timeout_cycles, int32_t* status) {
  __shared__ p2pfl_reduce_table s_red;
  if (threadIdx.x == 0) s_red = t;
  __syncthreads();
  reduce_phase(params, params_bf16, n, tile, seq, s_red, abort_flag, timeout_cycles, status, blockIdx.x, gridDim.x);
}

// G1+G2: one launch.  Every CTA first pushes its tiles (no waiting), then
// reduces the same tile indices; peers run the same schedule, so tile i of
// every contributor is in flight while tile i-1 is being accumulated.
__global__ void __launch_bounds__(kThreads, 2)
gossip_fused_kernel(float* __restrict__ params, __nv_bfloat16* __restrict__ params_bf16, uint64_t n, uint32_t tile,
                    uint32_t seq, const __grid_constant__ p2pfl_push_table push,
                    const __grid_constant__ p2pfl_reduce_table red, const uint32_t* __restrict__ abort_flag,
                    uint64_t timeout_cycles, int32_t* status, float* __restrict__ snapshot) {
  __shared__ p2pfl_push_table s_push;
  __shared__ p2pfl_reduce_table s_red;
  if (threadIdx.x == 0) {
    s_push = push;
    s_red = red;
  }
  __syncthreads();
  fused_phase(params, params_bf16, n, tile, seq, s_push, s_red, abort_flag, timeout_cycles, status, blockIdx.x,
              gridDim.x, snapshot);
}

// Several federated nodes that share this GPU exchange in ONE cooperative launch
// (blockIdx.y = local member): their tables may name inboxes on this GPU (the other
// local members) and on peer GPUs alike.  Waits between local members are safe because
// every CTA of a cooperative launch is resident; waits on peer GPUs are the ordinary
// cross-device case.  (B200_PROFILING.md: kernels that wait on one another must never be
// separate launches on one device.)
struct GroupArgs {
  const p2pfl_group_member* members;  // device array [n_members]
  const p2pfl_push_table* push;       // device array [n_members]
  const p2pfl_reduce_table* reduce;   // device array [n_members]
};

__global__ void __launch_bounds__(kThreads, 2)
gossip_fused_group_kernel(GroupArgs a, uint64_t n, uint32_t tile, uint32_t seq, uint64_t timeout_cycles) {
  const int r = blockIdx.y;
  __shared__ p2pfl_push_table s_push;
  __shared__ p2pfl_reduce_table s_red;
  __shared__ p2pfl_group_member s_me;
  if (threadIdx.x == 0) {
    s_push = a.push[r];
    s_red = a.reduce[r];
    s_me = a.members[r];
  }
  __syncthreads();
  fused_phase(s_me.params, reinterpret_cast<__nv_bfloat16*>(s_me.params_bf16), n, tile, seq, s_push, s_red,
              s_me.abort_flag, timeout_cycles, s_me.status, blockIdx.x, gridDim.x, s_me.snapshot);
}

// Test harness: R logical ranks in ONE cooperative launch on one GPU
// (blockIdx.y = rank).  Co-residency of all CTAs is what makes the cross-rank
// flag waits safe on a single device.
struct EmuArgs {
  float* params[P2PFL_MAX_PEERS];
  const p2pfl_push_table* push;     // device array [R]
  const p2pfl_reduce_table* reduce; // device array [R]
};

__global__ void __launch_bounds__(kThreads, 2)
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
  fused_phase(a.params[r], nullptr, n, tile, seq, s_push, s_red, nullptr, timeout_cycles, status + 4 * r, blockIdx.x,
              gridDim.x);
}

int check_common(uint64_t n, uint32_t tile) {
  P2PFL_REQUIRE(n > 0 && (n % 8) == 0, "gossip kernels need n to be a positive multiple of 8 elements");
  P2PFL_REQUIRE(tile >= 4096 && (tile % 4096) == 0, "tile must be a multiple of 4096 elements");
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
  return p2pfl_gossip_fused_snap(device, stream, params, params_bf16, n, tile, seq, push, reduce, abort_flag,
                                 timeout_cycles, status, nullptr);
}

int p2pfl_gossip_fused_snap(int device, void* stream, float* params, void* params_bf16, uint64_t n, uint32_t tile,
                            uint32_t seq, const p2pfl_push_table* push, const p2pfl_reduce_table* reduce,
                            const uint32_t* abort_flag, uint64_t timeout_cycles, int32_t* status, float* snapshot) {
  if (int rc = check_common(n, tile)) return rc;
  P2PFL_REQUIRE(push && reduce, "null table");
  P2PFL_REQUIRE(push->n_peers <= P2PFL_MAX_PEERS && reduce->n_slots <= P2PFL_MAX_PEERS, "too many peers");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  // the kernel spins on remote flags: keep the grid to ONE resident wave
  uint64_t tiles = (n + tile - 1) / tile;
  int per_sm = 0;
  P2PFL_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gossip_fused_kernel, kThreads, 0));
  if (per_sm > 2) per_sm = 2;
  if (per_sm < 1) per_sm = 1;
  uint64_t cap = (uint64_t)sm_count(device) * per_sm;
  int grid = (int)(tiles < cap ? tiles : cap);
  gossip_fused_kernel<<<grid, kThreads, 0, as_stream(stream)>>>(params, (__nv_bfloat16*)params_bf16, n, tile, seq,
                                                               *push, *reduce, abort_flag, timeout_cycles, status,
                                                               snapshot);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_gossip_fused_group(int device, void* stream, int n_members, const p2pfl_group_member* members_dev,
                             uint64_t n, uint32_t tile, uint32_t seq, const p2pfl_push_table* push_tables_dev,
                             const p2pfl_reduce_table* reduce_tables_dev, uint64_t timeout_cycles) {
  if (int rc = check_common(n, tile)) return rc;
  P2PFL_REQUIRE(n_members >= 1 && n_members <= P2PFL_MAX_PEERS, "a group exchange takes 1..16 local members");
  P2PFL_REQUIRE(members_dev && push_tables_dev && reduce_tables_dev, "null table");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  GroupArgs a;
  a.members = members_dev;
  a.push = push_tables_dev;
  a.reduce = reduce_tables_dev;
  int per_sm = 0;
  P2PFL_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gossip_fused_group_kernel, kThreads, 0));
  if (per_sm > 2) per_sm = 2;
  const int capacity = per_sm * sm_count(device);
  int gx = capacity / n_members;
  const uint64_t tiles = (n + tile - 1) / tile;
  if ((uint64_t)gx > tiles) gx = (int)tiles;
  P2PFL_REQUIRE(gx >= 1, "not enough resident CTAs for the group exchange");
  dim3 grid(gx, n_members, 1), block(kThreads, 1, 1);
  void* args[] = {&a, &n, &tile, &seq, &timeout_cycles};
  P2PFL_CUDA_OK(cudaLaunchCooperativeKernel((void*)gossip_fused_group_kernel, grid, block, args, 0, as_stream(stream)));
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

// ===========================================================================
// Allocation-free exchange for nodes that share one GPU.
//
// Some hosts that embed this library cannot obtain device memory when an
// exchange is due: a simulator that packs many federated nodes on one GPU and
// hands the whole device to its own pool at start-up has nothing left to ask
// cudaMalloc for, and a latency-sensitive round loop cannot afford the implicit
// device-wide synchronisation an allocation (or a free) performs.  For them the
// inbox the exchange needs lives in the library image itself.
//
// The kernel is the one behind p2pfl_gossip_fused_emulated; only the source of
// its scratch differs.  Because the ranks are co-located, every rank publishes
// its bf16 model ONCE into wire[rank] (instead of once per receiver) and all
// other ranks read it from there.
//
// Limits: 2..8 ranks, at most 262144 elements per model (the 235146-parameter
// MLP fits), bf16 wire.  The scratch is 4200192 bytes; every GPU that loads
// the library carries it until the process exits.
namespace {

constexpr int kStaticMaxRanks = 8;
constexpr uint64_t kStaticMaxElems = 262144;
constexpr uint32_t kStaticTile = 16384;
constexpr uint64_t kStaticTilesPerRank = kStaticMaxElems / kStaticTile;

constexpr uint64_t kOffWire = 0;
constexpr uint64_t kOffFlags = kOffWire + kStaticMaxRanks * kStaticMaxElems * 2;
constexpr uint64_t kOffStatus = kOffFlags + kStaticMaxRanks * kStaticTilesPerRank * 4;
constexpr uint64_t kOffPush = kOffStatus + 256;
constexpr uint64_t kOffReduce = kOffPush + 2304;  // 8 push tables of 264 bytes, rounded up
constexpr uint64_t kStaticScratchBytes = kOffReduce + 2816;  // 8 reduce tables of 344 bytes, rounded up
static_assert(sizeof(p2pfl_push_table) * kStaticMaxRanks <= 2304, "push tables outgrew their region");
static_assert(sizeof(p2pfl_reduce_table) * kStaticMaxRanks <= 2816, "reduce tables outgrew their region");
static_assert(kStaticScratchBytes == 4200192, "documented scratch size is stale");

// In host code this name is only a stand-in; the device address comes from
// cudaGetSymbolAddress (see static_scratch_for).
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

struct StaticSetup {
  float* params[kStaticMaxRanks];
  float weight[kStaticMaxRanks];
  int32_t n_ranks;
};

// One thread per rank writes that rank's push and reduce table into the scratch.
__global__ void static_setup_kernel(StaticSetup s, unsigned char* scratch) {
  const int r = threadIdx.x;
  if (r >= s.n_ranks) return;
  __nv_bfloat16* wire = reinterpret_cast<__nv_bfloat16*>(scratch + kOffWire);
  uint32_t* flags = reinterpret_cast<uint32_t*>(scratch + kOffFlags);
  p2pfl_push_table* push = reinterpret_cast<p2pfl_push_table*>(scratch + kOffPush) + r;
  p2pfl_reduce_table* red = reinterpret_cast<p2pfl_reduce_table*>(scratch + kOffReduce) + r;
  push->inbox[0] = wire + (uint64_t)r * kStaticMaxElems;
  push->flags[0] = flags + (uint64_t)r * kStaticTilesPerRank;
  push->n_peers = 1;
  push->wire_bf16 = 1;
  int k = 0;
  for (int p = 0; p < s.n_ranks; ++p) {
    if (p == r) continue;
    red->slot[k] = wire + (uint64_t)p * kStaticMaxElems;
    red->flags[k] = flags + (uint64_t)p * kStaticTilesPerRank;
    red->weight[k] = s.weight[p];
    ++k;
  }
  red->n_slots = k;
  red->self_pos = r;  // ranks are summed in ascending order, the local model at its own rank
  red->self_weight = s.weight[r];
  red->wire_bf16 = 1;
  red->round_self = 1;
  red->pad_ = 0;
}

// Device address of the scratch on `device`, resolved once per device.
int static_scratch_for(int device, unsigned char** out) {
  static std::mutex mu;
  static unsigned char* ptrs[64] = {nullptr};
  P2PFL_REQUIRE(device >= 0 && device < 64, "device ordinal out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[device] == nullptr) {
    unsigned char* p = nullptr;
    P2PFL_CUDA_OK(cudaSetDevice(device));
    P2PFL_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch));
    P2PFL_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    ptrs[device] = p;
  }
  *out = ptrs[device];
  return 0;
}

}  // namespace

extern "C" int p2pfl_gossip_exchange_static(int device, void* stream, int n_ranks, float* const* params,
                                            const float* weights, uint64_t n, uint64_t timeout_cycles,
                                            int32_t* status_out) {
  P2PFL_REQUIRE(params != nullptr && weights != nullptr, "null argument");
  if (n_ranks < 2 || n_ranks > kStaticMaxRanks) {
    set_error("p2pfl_gossip_exchange_static handles 2..%d co-located ranks, got %d", kStaticMaxRanks, n_ranks);
    return -3;
  }
  if (n == 0 || n > kStaticMaxElems || (n % 8) != 0) {
    set_error("p2pfl_gossip_exchange_static handles models of at most %llu elements (a multiple of 8), got %llu; "
              "use p2pfl_gossip_fused_emulated with caller-owned inboxes for larger models",
              (unsigned long long)kStaticMaxElems, (unsigned long long)n);
    return -3;
  }
  P2PFL_CUDA_OK(cudaSetDevice(device));
  unsigned char* scratch = nullptr;
  if (int rc = static_scratch_for(device, &scratch)) return rc;
  cudaStream_t s = as_stream(stream);

  // The scratch still holds the previous call's flags and status: arrival flags and the
  // status words must start at zero (the wire data and the tables are overwritten below).
  P2PFL_CUDA_OK(cudaMemsetAsync(scratch + kOffFlags, 0, kOffPush - kOffFlags, s));

  StaticSetup setup;
  EmuArgs a;
  for (int r = 0; r < kStaticMaxRanks; ++r) {
    setup.params[r] = r < n_ranks ? params[r] : nullptr;
    setup.weight[r] = r < n_ranks ? weights[r] : 0.f;
    a.params[r] = setup.params[r];
    if (r < n_ranks) P2PFL_REQUIRE(params[r] != nullptr && ((uintptr_t)params[r] & 15) == 0, "params must be 16-byte aligned");
  }
  for (int r = kStaticMaxRanks; r < P2PFL_MAX_PEERS; ++r) a.params[r] = nullptr;
  setup.n_ranks = n_ranks;
  static_setup_kernel<<<1, 32, 0, s>>>(setup, scratch);
  P2PFL_LAUNCH_OK();

  a.push = reinterpret_cast<const p2pfl_push_table*>(scratch + kOffPush);
  a.reduce = reinterpret_cast<const p2pfl_reduce_table*>(scratch + kOffReduce);
  int32_t* status = reinterpret_cast<int32_t*>(scratch + kOffStatus);
  uint32_t tile = kStaticTile, seq = 1;
  int per_sm = 0;
  P2PFL_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gossip_fused_emulated_kernel, kThreads, 0));
  int gx = per_sm * sm_count(device) / n_ranks;
  const uint64_t tiles = (n + tile - 1) / tile;
  if ((uint64_t)gx > tiles) gx = (int)tiles;
  P2PFL_REQUIRE(gx >= 1, "not enough resident CTAs for the co-located exchange");
  dim3 grid(gx, n_ranks, 1), block(kThreads, 1, 1);
  void* args[] = {&a, &n, &tile, &seq, &timeout_cycles, &status};
  P2PFL_CUDA_OK(cudaLaunchCooperativeKernel((void*)gossip_fused_emulated_kernel, grid, block, args, 0, s));
  if (status_out) P2PFL_CUDA_OK(cudaMemcpyAsync(status_out, status, sizeof(int32_t) * 4 * n_ranks, cudaMemcpyDeviceToDevice, s));
  return 0;
}
