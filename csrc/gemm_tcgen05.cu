// D[M,N] = A[M,K] . B[N,K]^T on Blackwell 5th-gen tensor cores.
//
// reference sites replaced: mlp.py:63-67 / cnn.py:84-85 (nn.Linear forward,
// cuBLAS sgemm + bias + ReLU) and their autograd dgrad / wgrad GEMMs
// (SURVEY.md K1, K2, K4, K8).
//
// Structure (one 128 x BN output tile per CTA, 6 warps):
//   warp 0      TMA producer : cp.async.bulk.tensor (128B swizzle) global->smem,
//                              mbarrier ring sized to the K loop, complete_tx
//   warp 1      MMA issuer   : one elected lane issues tcgen05.mma (kind::f16,
//                              bf16 x bf16 -> fp32) with smem descriptors;
//                              accumulator lives in TMEM; tcgen05.commit frees
//                              smem stages and finally signals the epilogue
//   warps 2..5  epilogue     : tcgen05.ld TMEM->registers, fused
//                              alpha / bias / ReLU / ReLU-mask / row-sum, then
//                              up to four output layouts (fp32|bf16, direct|transposed)
//
// Operands are K-major by default; an operand may also be MN-major (stored [K, MN]), which lets
// the backward GEMMs of the convolutions read dZ, the column matrix and the weights in place.
// Kernels in this file:
//   gemm_tn_kernel             one tile per CTA, up to 4 problems per launch, optional split-K
//   gemm_tn_persistent_kernel  one CTA per SM walks the tiles, two TMEM accumulators
//   gemm_chain_kernel          a list of dependent links walked by one CTA or by ONE thread-block
//                              cluster (tiles dealt to its CTAs, barrier.cluster between links)
#include <cstdlib>
#include <cuda.h>
#include <string.h>

#include "common.cuh"
#include "tcgen05_ptx.cuh"

using namespace p2pfl;
using namespace p2pfl::tc;

namespace {

constexpr int kMaxStages = 10;  // barriers are sized for this; the ring depth is chosen per problem
constexpr int kThreadsGemm = 192;
constexpr uint32_t kTileABytes = BM * BK * 2;

struct Epilogue {
  const float* bias;
  const __nv_bfloat16* mask;
  float* out_f32;
  __nv_bfloat16* out_bf16;
  float* out_f32_t;
  __nv_bfloat16* out_bf16_t;
  float* rowsum;
  float alpha;
  int32_t ld_mask, ld_f32, ld_bf16, ld_f32_t, ld_bf16_t;
  int32_t bias_per_row, relu;
  int32_t atomic;  // split-K: out_f32 is accumulated with red.global.add (caller zero-fills it)
};

// Drain one 128 x bn accumulator tile: TMEM -> registers -> fused epilogue -> global.
// Called by the four epilogue warps; `q` is the warp's TMEM lane quarter.
__device__ __forceinline__ void epilogue_tile(const Epilogue& ep_mem, int M, int N, int m0, int n0, uint32_t bn,
                                              uint32_t tmem_base, int q, int lane, bool single_n_tile
#ifdef P2PFL_BRINGUP
                                              , long long* dbg = nullptr
#endif
) {
  // by value: read through the reference (shared or constant memory), every field would be
  // reloaded after each global store - the stores might alias it as far as the compiler knows
  const Epilogue ep = ep_mem;
  const int m = m0 + q * 32 + lane;
#ifdef P2PFL_BRINGUP
#define EPI_STAMP(slot) \
  if (dbg && lane == 0) dbg[slot] = clock64()
#else
#define EPI_STAMP(slot)
#endif
  EPI_STAMP(0);
  const float bias_row = (ep.bias && ep.bias_per_row && m < M) ? ep.bias[m] : 0.f;
#ifdef P2PFL_BRINGUP
  if (dbg && bias_row == 123456.f) dbg[7] = 1;  // (forces the load to complete before the next stamp)
#endif
  EPI_STAMP(1);
#pragma unroll 1
  for (int c = 0; c < (int)bn / 32; ++c) {
    const int nb = n0 + c * 32;
    if (nb >= N) break;
    uint32_t r[32];
    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(c * 32), r);
    EPI_STAMP(2);
    // the option tests stay OUTSIDE the element loops: inside, each becomes a branch per element
    float x[32];
    if (ep.bias && !ep.bias_per_row) {
#pragma unroll
      for (int j = 0; j < 32; ++j) x[j] = (nb + j < N) ? __ldg(ep.bias + nb + j) : 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) x[j] = fmaf(__uint_as_float(r[j]), ep.alpha, x[j]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) x[j] = fmaf(__uint_as_float(r[j]), ep.alpha, bias_row);
    }
    if (ep.relu) {
#pragma unroll
      for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.f);
    }
    if (ep.mask && m < M) {
      const __nv_bfloat16* mk = ep.mask + (size_t)m * ep.ld_mask + nb;
      if ((N - nb) >= 32 && ((reinterpret_cast<uintptr_t>(mk) & 15) == 0)) {
        // the row's 32 mask values as four 16-byte loads, all in flight together
        uint4 mv[4];
#pragma unroll
        for (int v = 0; v < 4; ++v) mv[v] = *reinterpret_cast<const uint4*>(mk + v * 8);
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          const uint32_t w[4] = {mv[v].x, mv[v].y, mv[v].z, mv[v].w};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (!(__uint_as_float(w[k] << 16) > 0.f)) x[v * 8 + 2 * k] = 0.f;
            if (!(__uint_as_float(w[k] & 0xFFFF0000u) > 0.f)) x[v * 8 + 2 * k + 1] = 0.f;
          }
        }
      } else {
        float mf[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) mf[j] = (nb + j < N) ? __bfloat162float(mk[j]) : 1.f;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (!(mf[j] > 0.f)) x[j] = 0.f;
      }
    }
    const int valid = (N - nb) < 32 ? (N - nb) : 32;
#ifdef P2PFL_BRINGUP
    if (dbg && x[0] == 123456.f) dbg[7] = 2;
#endif
    EPI_STAMP(3);
    if (m < M) {
      if (ep.rowsum) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j < valid) s += x[j];
        // one N tile of <= 32 columns owns the whole row: plain store, no zero-init needed
        if (single_n_tile && N <= 32) ep.rowsum[m] = s;
        else atomicAdd(ep.rowsum + m, s);
      }
      // direct ([M, ld]) outputs: each thread writes its row's 32 contiguous values
      if (ep.out_f32 && ep.atomic) {
        // L2 retires a fixed number of reductions per clock whatever their width: use 16-byte ones
        float* o = ep.out_f32 + (size_t)m * ep.ld_f32 + nb;
        if (valid == 32 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + j), "f"(x[j]), "f"(x[j + 1]),
                         "f"(x[j + 2]), "f"(x[j + 3])
                         : "memory");
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < valid) atomicAdd(o + j, x[j]);
        }
      } else if (ep.out_f32) {
        float* o = ep.out_f32 + (size_t)m * ep.ld_f32 + nb;
        if (valid == 32 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(x[j], x[j + 1], x[j + 2], x[j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < valid) o[j] = x[j];
        }
      }
      if (ep.out_bf16) {
        __nv_bfloat16* o = ep.out_bf16 + (size_t)m * ep.ld_bf16 + nb;
        if (valid == 32 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            uint4 u;
            u.x = pack_bf16x2(x[j], x[j + 1]);
            u.y = pack_bf16x2(x[j + 2], x[j + 3]);
            u.z = pack_bf16x2(x[j + 4], x[j + 5]);
            u.w = pack_bf16x2(x[j + 6], x[j + 7]);
            *reinterpret_cast<uint4*>(o + j) = u;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < valid) o[j] = __float2bfloat16_rn(x[j]);
        }
      }
      EPI_STAMP(4);
      // transposed outputs: consecutive lanes hold consecutive m -> coalesced per column
      if (ep.out_f32_t) {
        float* o = ep.out_f32_t + (size_t)nb * ep.ld_f32_t + m;
        if (valid == 32) {
#pragma unroll
          for (int j = 0; j < 32; ++j) o[(size_t)j * ep.ld_f32_t] = x[j];
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < valid) o[(size_t)j * ep.ld_f32_t] = x[j];
        }
      }
      if (ep.out_bf16_t) {
        __nv_bfloat16* o = ep.out_bf16_t + (size_t)nb * ep.ld_bf16_t + m;
        if (valid == 32) {
#pragma unroll
          for (int j = 0; j < 32; ++j) o[(size_t)j * ep.ld_bf16_t] = __float2bfloat16_rn(x[j]);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < valid) o[(size_t)j * ep.ld_bf16_t] = __float2bfloat16_rn(x[j]);
        }
      }
      EPI_STAMP(5);
    }
  }
#undef EPI_STAMP
}

constexpr int kMaxGroup = 4;
constexpr int kMaxBN = 256;
constexpr uint32_t kRingBytes = 4 * (kTileABytes + kMaxBN * BK * 2);  // 4 stages of the widest tile = 192 KB

// Up to four independent GEMMs per launch (blockIdx.z picks one); each has its own
// tensor maps, epilogue and N-tile width bn in {32, 64, 128, 256}.
struct GroupedArgs {
  CUtensorMap ma[kMaxGroup];
  CUtensorMap mb[kMaxGroup];
  Epilogue ep[kMaxGroup];
  int32_t M[kMaxGroup], N[kMaxGroup], K[kMaxGroup], bn[kMaxGroup];
  int32_t split[kMaxGroup];  // K slices per output tile (>= 1); slices accumulate atomically
  int32_t mn[kMaxGroup];     // bit 0: A is MN-major ([K, lda]), bit 1: B is MN-major ([K, ldb])
  int32_t ring_bytes;        // dynamic shared memory given to the stage ring of the one-shot kernel
  int32_t count;
};

// ------------------------------------------------------------------- kernel
__global__ void __launch_bounds__(kThreadsGemm, 1)
gemm_tn_kernel(const __grid_constant__ GroupedArgs g) {
  const int prob = blockIdx.z;
  const int M = g.M[prob], N = g.N[prob], K = g.K[prob];
  const uint32_t bn = (uint32_t)g.bn[prob];
  const int split = g.split[prob];
  const uint32_t mn = (uint32_t)g.mn[prob];
  const int m0 = (int)(blockIdx.x / split) * BM;
  const int n0 = blockIdx.y * (int)bn;
  if (m0 >= M || n0 >= N) return;  // grid is the union of the problems' tile grids
  // this CTA's slice of the K loop
  const int kb_all = (K + BK - 1) / BK;
  const int kb_per = (kb_all + split - 1) / split;
  const int kb0 = (int)(blockIdx.x % split) * kb_per;
  if (kb0 >= kb_all) return;
  const CUtensorMap* tma_a = &g.ma[prob];
  const CUtensorMap* tma_b = &g.mb[prob];
  const Epilogue& ep = g.ep[prob];
  const uint32_t stage_tx = kTileABytes + bn * BK * 2;
  // narrow tiles get a deeper ring out of the same shared memory: the small-batch GEMMs of
  // the train step are latency bound, and more loads in flight shorten their K loop
  const uint32_t stage_bytes = stage_tx;
  // the launch sizes the ring for the longest K loop of its problems: short loops take little
  // shared memory, so several CTAs share an SM and hide each other's load and drain latency
  int n_stages = (int)((uint32_t)g.ring_bytes / stage_bytes);
  n_stages = n_stages > kMaxStages ? kMaxStages : n_stages;

  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles must start on 1024-byte boundaries
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t full_bar[kMaxStages];
  __shared__ __align__(8) uint64_t empty_bar[kMaxStages];
  __shared__ __align__(8) uint64_t tmem_full_bar;
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_kb = (kb_all - kb0) < kb_per ? (kb_all - kb0) : kb_per;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tma_a)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tma_b)) : "memory");
    for (int s = 0; s < n_stages; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    mbar_init(smem_u32(&tmem_full_bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_smem), bn);
  pdl_launch_dependents();  // let the next kernel of the chain start its own prologue
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();  // everything above overlapped the predecessor; its outputs are visible from here on

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % n_stages;
        const uint32_t phase = (uint32_t)(kb / n_stages) & 1u;
        mbar_wait(smem_u32(&empty_bar[s]), phase ^ 1u);
        const uint32_t bar = smem_u32(&full_bar[s]);
        mbar_expect_tx(bar, stage_tx);
        const uint32_t a_dst = smem_u32(smem + (size_t)s * stage_bytes);
        const int k_crd = (kb0 + kb) * BK;
        if (mn & 1u) {
          for (int j = 0; j < BM / 64; ++j) tma_load_2d(tma_a, bar, a_dst + j * kMnBoxBytes, m0 + j * 64, k_crd);
        } else {
          tma_load_2d(tma_a, bar, a_dst, k_crd, m0);
        }
        if (mn & 2u) {
          for (int j = 0; j < (int)bn / 64; ++j)
            tma_load_2d(tma_b, bar, a_dst + kTileABytes + j * kMnBoxBytes, n0 + j * 64, k_crd);
        } else {
          tma_load_2d(tma_b, bar, a_dst + kTileABytes, k_crd, n0);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t idesc = make_idesc(bn, mn);
    // bytes >> 4 per 16-element K step: 32 B inside a K-major swizzle row, two 1024 B atoms MN-major
    const uint64_t a_step = (mn & 1u) ? (2048 >> 4) : 2;
    const uint64_t b_step = (mn & 2u) ? (2048 >> 4) : 2;
    for (int kb = 0; kb < num_kb; ++kb) {
      const int s = kb % n_stages;
      const uint32_t phase = (uint32_t)(kb / n_stages) & 1u;
      mbar_wait(smem_u32(&full_bar[s]), phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t a_addr = smem_u32(smem + (size_t)s * stage_bytes);
        const uint64_t da = (mn & 1u) ? make_smem_desc_mn(a_addr) : make_smem_desc(a_addr);
        const uint64_t db = (mn & 2u) ? make_smem_desc_mn(a_addr + kTileABytes) : make_smem_desc(a_addr + kTileABytes);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          // advance 16 bf16 along K: a_step / b_step in the (addr >> 4) field
          umma_bf16(tmem_base, da + a_step * (uint64_t)k, db + b_step * (uint64_t)k, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit(smem_u32(&empty_bar[s]));  // smem stage reusable once these MMAs retire
        if (kb == num_kb - 1) umma_commit(smem_u32(&tmem_full_bar));
      }
      __syncwarp();
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    // a warp may only touch the TMEM lane quarter given by (warp id % 4)
    const int q = warp & 3;
    mbar_wait(smem_u32(&tmem_full_bar), 0);
    tc_fence_after();
    epilogue_tile(ep, M, N, m0, n0, bn, tmem_base, q, lane, gridDim.y == 1 || (int)bn >= N);
    tc_fence_before();
  }

  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, bn);
  }
}

// ------------------------------------------------------ persistent variant
// Large problems (many more tiles than SMs): one CTA per SM walks its share of the
// output tiles.  Barriers, TMEM and the smem ring are set up once; TWO accumulators in
// TMEM (2 x bn columns) let the MMA warp start tile i+1 while the epilogue warps are
// still draining tile i, so the tensor pipe does not idle during epilogues.
// Tiles are visited M-fastest inside an N panel, which keeps the B panel in L2.
__global__ void __launch_bounds__(kThreadsGemm, 1)
gemm_tn_persistent_kernel(const __grid_constant__ GroupedArgs g) {
  const int M = g.M[0], N = g.N[0], K = g.K[0];
  const uint32_t bn = (uint32_t)g.bn[0];
  const uint32_t mn = (uint32_t)g.mn[0];
  const CUtensorMap* tma_a = &g.ma[0];
  const CUtensorMap* tma_b = &g.mb[0];
  const Epilogue& ep = g.ep[0];
  const uint32_t stage_bytes = kTileABytes + bn * BK * 2;
  int n_stages = (int)(kRingBytes / stage_bytes);
  n_stages = n_stages > kMaxStages ? kMaxStages : n_stages;
  const int tiles_m = (M + BM - 1) / BM, tiles_n = (N + (int)bn - 1) / (int)bn;
  const int n_tiles = tiles_m * tiles_n;
  const int num_kb = (K + BK - 1) / BK;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t full_bar[kMaxStages];
  __shared__ __align__(8) uint64_t empty_bar[kMaxStages];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tma_a)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tma_b)) : "memory");
    for (int s = 0; s < n_stages; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&tmem_full_bar[a]), 1);
      mbar_init(smem_u32(&tmem_empty_bar[a]), 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_smem), 2 * bn);
  pdl_launch_dependents();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();

  if (warp == 0) {
    if (elect_one()) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const int m0 = (t % tiles_m) * BM, n0 = (t / tiles_m) * (int)bn;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const uint32_t s = it % n_stages, ph = (it / n_stages) & 1u;
          mbar_wait(smem_u32(&empty_bar[s]), ph ^ 1u);
          const uint32_t bar = smem_u32(&full_bar[s]);
          mbar_expect_tx(bar, stage_bytes);
          const uint32_t dst = smem_u32(smem + (size_t)s * stage_bytes);
          if (mn & 1u) {
            for (int j = 0; j < BM / 64; ++j) tma_load_2d(tma_a, bar, dst + j * kMnBoxBytes, m0 + j * 64, kb * BK);
          } else {
            tma_load_2d(tma_a, bar, dst, kb * BK, m0);
          }
          if (mn & 2u) {
            for (int j = 0; j < (int)bn / 64; ++j)
              tma_load_2d(tma_b, bar, dst + kTileABytes + j * kMnBoxBytes, n0 + j * 64, kb * BK);
          } else {
            tma_load_2d(tma_b, bar, dst + kTileABytes, kb * BK, n0);
          }
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = make_idesc(bn, mn);
    const uint64_t a_step = (mn & 1u) ? (2048 >> 4) : 2;
    const uint64_t b_step = (mn & 2u) ? (2048 >> 4) : 2;
    uint32_t it = 0, tile_it = 0;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++tile_it) {
      const uint32_t acc = tile_it & 1u, aph = (tile_it >> 1) & 1u;
      mbar_wait(smem_u32(&tmem_empty_bar[acc]), aph ^ 1u);
      tc_fence_after();
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        const uint32_t s = it % n_stages, ph = (it / n_stages) & 1u;
        mbar_wait(smem_u32(&full_bar[s]), ph);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a_addr = smem_u32(smem + (size_t)s * stage_bytes);
          const uint64_t da = (mn & 1u) ? make_smem_desc_mn(a_addr) : make_smem_desc(a_addr);
          const uint64_t db = (mn & 2u) ? make_smem_desc_mn(a_addr + kTileABytes) : make_smem_desc(a_addr + kTileABytes);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            umma_bf16(tmem_base + acc * bn, da + a_step * (uint64_t)k, db + b_step * (uint64_t)k, idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit(smem_u32(&empty_bar[s]));
          if (kb == num_kb - 1) umma_commit(smem_u32(&tmem_full_bar[acc]));
        }
        __syncwarp();
      }
    }
  } else {
    const int q = warp & 3;
    uint32_t tile_it = 0;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++tile_it) {
      const uint32_t acc = tile_it & 1u, aph = (tile_it >> 1) & 1u;
      const int m0 = (t % tiles_m) * BM, n0 = (t / tiles_m) * (int)bn;
      mbar_wait(smem_u32(&tmem_full_bar[acc]), aph);
      tc_fence_after();
      epilogue_tile(ep, M, N, m0, n0, bn, tmem_base + acc * bn, q, lane, tiles_n == 1);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&tmem_empty_bar[acc]));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * bn);
  }
}

// ------------------------------------------------------------- host: TMA maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 [rows, ld] row-major, logical width `cols` (<= ld); box = [64 x box_rows], 128B swizzle
int make_map(CUtensorMap* map, const void* base, int rows, int cols, int ld, int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  P2PFL_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  P2PFL_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "GEMM operands must be 16-byte aligned");
  P2PFL_REQUIRE(((size_t)ld * 2) % 16 == 0, "GEMM operand leading dimensions must be multiples of 8 elements");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%d cols=%d ld=%d box_rows=%d)", (int)r, rows, cols,
              ld, box_rows);
    return -2;
  }
  return 0;
}

// bf16 [k_rows, ld] row-major holding an MN-major operand: logical width `mn_cols`; box = 64 MN
// elements (one 128-byte swizzle row) x BK k-rows
int make_map_mn(CUtensorMap* map, const void* base, int k_rows, int mn_cols, int ld) {
  EncodeTiledFn enc = get_encode_fn();
  P2PFL_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  P2PFL_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "GEMM operands must be 16-byte aligned");
  P2PFL_REQUIRE(((size_t)ld * 2) % 16 == 0, "GEMM operand leading dimensions must be multiples of 8 elements");
  cuuint64_t gdim[2] = {(cuuint64_t)mn_cols, (cuuint64_t)k_rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64u, (cuuint32_t)BK};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (MN-major, k_rows=%d cols=%d ld=%d)", (int)r, k_rows,
              mn_cols, ld);
    return -2;
  }
  return 0;
}

// N-tile width: the widest that fits, but narrow it while the grid would leave most SMs
// idle - small-K problems (wgrad of a small batch) are bound by the epilogue stores, which
// spread over more SMs with narrower tiles.
int pick_bn(int M, int N) {
  int bn = N <= 32 ? 32 : N <= 64 ? 64 : N <= 128 ? 128 : 256;
  const int tiles_m = (M + BM - 1) / BM;
  while (bn > 32 && tiles_m * ((N + bn - 1) / bn) < 74) bn >>= 1;
  return bn;
}

int fill_problem(GroupedArgs& g, int idx, const p2pfl_gemm_args& a) {
  P2PFL_REQUIRE(a.A && a.B && a.M > 0 && a.N > 0 && a.K > 0, "bad GEMM shape or operands");
  P2PFL_REQUIRE(a.out_f32 || a.out_bf16 || a.out_f32_t || a.out_bf16_t || a.rowsum, "GEMM needs an output");
  const int mn = a.mn_major & 3;
  const int tiles_m = (a.M + BM - 1) / BM;
  const int kb_all = (a.K + BK - 1) / BK;
  int bn, split = 1;
  if (a.split_k != 0 && a.split_k != 1) {
    // split-K problems (weight gradients: small output, long reduction): widest tile, then
    // enough K slices to fill the machine, at least four k-blocks each
    bn = a.N <= 32 ? 32 : a.N <= 64 ? 64 : a.N <= 128 ? 128 : 256;
    const int tiles = tiles_m * ((a.N + bn - 1) / bn);
    // one wave of CTAs: every extra slice costs another pass of reductions over D
    split = a.split_k > 1 ? a.split_k : (148 + tiles - 1) / tiles;
    if (split > kb_all / 4) split = kb_all / 4;
    if (split < 1) split = 1;
  } else {
    bn = pick_bn(a.M, a.N);
  }
  if ((mn & 2) && bn < 64) bn = 64;  // an MN-major B tile is made of 64-column boxes
  if (split > 1) {
    P2PFL_REQUIRE(a.out_f32 && !a.out_bf16 && !a.out_f32_t && !a.out_bf16_t && !a.rowsum && !a.bias && !a.relu && !a.mask,
                  "a split-K GEMM accumulates into out_f32 only (zero-filled by the caller), without bias / ReLU / mask");
  }
  if (mn & 1) {
    if (int rc = make_map_mn(&g.ma[idx], a.A, a.K, a.M, a.lda)) return rc;
  } else {
    if (int rc = make_map(&g.ma[idx], a.A, a.M, a.K, a.lda, BM)) return rc;
  }
  if (mn & 2) {
    if (int rc = make_map_mn(&g.mb[idx], a.B, a.K, a.N, a.ldb)) return rc;
  } else {
    if (int rc = make_map(&g.mb[idx], a.B, a.N, a.K, a.ldb, bn)) return rc;
  }
  g.split[idx] = split;
  g.mn[idx] = mn;
  Epilogue& ep = g.ep[idx];
  ep.atomic = split > 1;
  ep.bias = a.bias;
  ep.mask = reinterpret_cast<const __nv_bfloat16*>(a.mask);
  ep.out_f32 = a.out_f32;
  ep.out_bf16 = reinterpret_cast<__nv_bfloat16*>(a.out_bf16);
  ep.out_f32_t = a.out_f32_t;
  ep.out_bf16_t = reinterpret_cast<__nv_bfloat16*>(a.out_bf16_t);
  ep.rowsum = a.rowsum;
  ep.alpha = a.alpha;
  ep.ld_mask = a.ld_mask;
  ep.ld_f32 = a.ld_f32;
  ep.ld_bf16 = a.ld_bf16;
  ep.ld_f32_t = a.ld_f32_t;
  ep.ld_bf16_t = a.ld_bf16_t;
  ep.bias_per_row = a.bias_per_row;
  ep.relu = a.relu;
  g.M[idx] = a.M;
  g.N[idx] = a.N;
  g.K[idx] = a.K;
  g.bn[idx] = bn;
  return 0;
}

int launch_group(int device, cudaStream_t stream, const p2pfl_gemm_args* args, int count) {
  P2PFL_REQUIRE(args != nullptr && count >= 1 && count <= kMaxGroup, "a grouped GEMM launch takes 1..4 problems");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  GroupedArgs g;
  memset(&g, 0, sizeof(g));
  g.count = count;
  int gx = 0, gy = 0;
  for (int i = 0; i < count; ++i) {
    if (int rc = fill_problem(g, i, args[i])) return rc;
    const int tx = (args[i].M + BM - 1) / BM * g.split[i], ty = (args[i].N + g.bn[i] - 1) / g.bn[i];
    gx = tx > gx ? tx : gx;
    gy = ty > gy ? ty : gy;
    const int stage = (int)kTileABytes + g.bn[i] * BK * 2;
    const int kb_cta = ((args[i].K + BK - 1) / BK + g.split[i] - 1) / g.split[i];
    int stages = (int)kRingBytes / stage;
    stages = stages > kMaxStages ? kMaxStages : stages;
    stages = stages > kb_cta ? kb_cta : stages;
    g.ring_bytes = g.ring_bytes > stages * stage ? g.ring_bytes : stages * stage;
  }
  const bool plain = g.split[0] == 1;
  constexpr size_t smem = (size_t)kRingBytes + 1024;
  static bool attr_set[64] = {false};
  if (device >= 0 && device < 64 && !attr_set[device]) {
    P2PFL_CUDA_OK(cudaFuncSetAttribute(gemm_tn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set[device] = true;
  }
  const int sms = sm_count(device);
  if (count == 1 && plain && gx * gy >= 2 * sms) {
    // many more tiles than SMs: persistent CTAs with double-buffered accumulators
    static bool pattr_set[64] = {false};
    if (device >= 0 && device < 64 && !pattr_set[device]) {
      P2PFL_CUDA_OK(cudaFuncSetAttribute(gemm_tn_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      pattr_set[device] = true;
    }
    P2PFL_CUDA_OK(launch_pdl(gemm_tn_persistent_kernel, dim3(sms), dim3(kThreadsGemm), smem, stream, g));
    return 0;
  }
  P2PFL_CUDA_OK(launch_pdl(gemm_tn_kernel, dim3(gx, gy, count), dim3(kThreadsGemm), (size_t)g.ring_bytes + 1024, stream, g));
  return 0;
}

}  // namespace

extern "C" int p2pfl_gemm_tn(int device, void* stream, const p2pfl_gemm_args* args) {
  return launch_group(device, as_stream(stream), args, 1);
}

extern "C" int p2pfl_gemm_tn_grouped(int device, void* stream, const p2pfl_gemm_args* args, int count) {
  return launch_group(device, as_stream(stream), args, count);
}

// ===========================================================================
// GEMM chain: the DEPENDENT part of a small-batch train step in one launch.
//
// The MLP step (mlp.py:57-80 + autograd) is a chain of tiny GEMMs in which every
// link needs the previous link's output: fwd1 -> fwd2 -> fwd3 -> softmax/CE ->
// dgrad3 -> dgrad2.  As separate launches each link pays a launch + prologue +
// drain (~5 us for ~0.1 us of math).  Here ONE CTA walks the whole list: barriers,
// TMEM and tensor maps are set up once, the TMA / MMA / epilogue warps keep their
// pipeline state across links, and a CTA-wide barrier with a generic->async proxy
// fence separates links (the next link's TMA loads read what this link's epilogue
// stored).  N (the batch) is tiled in 32 columns; two 32-column TMEM accumulators
// let the MMA of the next tile overlap the epilogue of the current one.
namespace {

constexpr int kChainMaxGemm = 8;
constexpr int kChainMaxPhases = 12;
// N-tile width.  The kernel carries a width per GEMM (chain_bn_for), but measured on B200 a
// 128-wide tile for the weight-gradient link was slower than 32-wide ones (the epilogue's first
// TMEM load stalled ~3 us behind the next tile's MMAs), so every link uses 32 columns.
constexpr uint32_t kChainBN = 32;
constexpr uint32_t kChainMaxBN = 32;
constexpr uint32_t kChainStageBytes = kTileABytes + kChainMaxBN * BK * 2;  // 20 KB
constexpr int kChainStages = 9;
__host__ __device__ __forceinline__ int chain_bn_for(int N) { return N >= 128 ? (int)kChainMaxBN : (int)kChainBN; }

// Everything but the tensor maps.  The kernel copies this block from the parameter (constant)
// space to shared memory once, all threads at once: read in place, every first touch of a
// constant line is a ~0.5 us miss and the links would pay them one after the other.
struct ChainScalars {
  Epilogue ep[kChainMaxGemm];
  int32_t M[kChainMaxGemm], N[kChainMaxGemm], K[kChainMaxGemm];
  int32_t tiles_n[kChainMaxGemm], tiles[kChainMaxGemm], num_kb[kChainMaxGemm];  // host-computed: no divisions per tile
  int32_t a_rows[kChainMaxGemm];  // rows of the A box: min(128, M rounded up to 8) - TMA time is per row, valid or not
  int32_t kind[kChainMaxPhases];  // 0 = GEMM(s), 1 = softmax cross-entropy
  int32_t idx[kChainMaxPhases];
  int32_t cnt[kChainMaxPhases];   // GEMM link: idx .. idx+cnt-1 are independent and share the link
  int32_t n_phases;
  int32_t pad_;
  p2pfl_ce_args ce;
#ifdef P2PFL_BRINGUP
  long long* stamps;  // [link][8] SM-clock stamps of CTA 0, see p2pfl_gemm_chain_set_stamps
#endif
};
static_assert(sizeof(ChainScalars) % 4 == 0, "copied word by word");

struct ChainArgs {
  CUtensorMap ma[kChainMaxGemm];
  CUtensorMap mb[kChainMaxGemm];
  ChainScalars sc;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_release_acquire() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// tile `t` of a link that spans GEMMs gi0 .. gi0+cnt-1 -> (GEMM, m0, n0, its k-block count)
struct ChainTile {
  int gi, m0, n0, num_kb, bn, a_rows;
};
__device__ __forceinline__ int chain_tiles_of(const ChainScalars& g, int gi) { return g.tiles[gi]; }
__device__ __forceinline__ ChainTile chain_locate(const ChainScalars& g, int gi0, int cnt, int t) {
  int gi = gi0;
  for (int j = 0; j + 1 < cnt; ++j) {
    const int n = chain_tiles_of(g, gi);
    if (t < n) break;
    t -= n;
    ++gi;
  }
  const int bn = chain_bn_for(g.N[gi]);
  const int tiles_n = g.tiles_n[gi];
  const int tm = tiles_n == 1 ? t : t / tiles_n;
  ChainTile ct;
  ct.gi = gi;
  ct.bn = bn;
  ct.m0 = tm * BM;
  ct.n0 = (t - tm * tiles_n) * bn;
  ct.num_kb = g.num_kb[gi];
  ct.a_rows = g.a_rows[gi];
  return ct;
}

// log-softmax + cross-entropy forward/backward by the 128 epilogue threads (one warp per row)
// `scratch`: idle shared memory (the stage ring between links) for B * (C + 1) words, or null.
// Staging the logits there costs ONE round trip to L2; read in place, every row pays three.
__device__ void chain_ce(const p2pfl_ce_args& a_mem, int ewarp, int lane, float* s_db, float* scratch, bool small
#ifdef P2PFL_BRINGUP
                         , long long* dbg = nullptr
#endif
) {
#ifdef P2PFL_BRINGUP
#define CE_STAMP(slot) \
  if (dbg && ewarp == 0 && lane == 0) dbg[slot] = clock64()
#else
#define CE_STAMP(slot)
#endif
  CE_STAMP(0);
  const p2pfl_ce_args a = a_mem;  // by value, for the same reason as in epilogue_tile
  const int et = ewarp * 32 + lane;
  const int C = a.C, B = a.B;
  for (int c = et; c < C; c += 128) s_db[c] = 0.f;
  int* s_lab = reinterpret_cast<int*>(scratch + (size_t)B * C);
  if (scratch && !small) {
    for (int i = et; i < B * C; i += 128) scratch[i] = __ldcg(a.logits + (size_t)(i / C) * a.ld_logits + (i % C));
    for (int i = et; i < B; i += 128) s_lab[i] = a.labels[i];
  }
  asm volatile("bar.sync 1, 128;" ::: "memory");
  const float inv_b = 1.f / (float)B;
  float wloss = 0.f;
  __nv_bfloat16* dl = reinterpret_cast<__nv_bfloat16*>(a.dlogits_bf16);
  __nv_bfloat16* dlt = reinterpret_cast<__nv_bfloat16*>(a.dlogits_t_bf16);
  if (scratch && small) {
    // few classes, at most 128 rows.  Three short phases, each with all the parallelism there is:
    //   (1) all 128 threads stage logits + labels in shared memory (one trip to L2),
    //   (2) one thread per row: max / sum-exp over its C classes -> log-sum-exp, loss,
    //   (3) all threads: one gradient element each (both orientations + a staged copy),
    // then C threads sum the staged gradients into the bias gradient.
    // rows padded to a multiple of 4 words (16-byte aligned rows; the padding holds -inf)
    const int CP = (C + 3) & ~3;
    float* xs = scratch;                                    // [B, CP]
    int* lab = reinterpret_cast<int*>(scratch + B * CP);    // [B]
    float* s_lse = scratch + B * CP + B;                    // [B]
    float* gs = s_lse + B;                                  // [B, C]
    // four loads in flight per thread before the first store: one trip to L2, not one per element
    const int lab_mine = et < B ? a.labels[et] : 0;
    for (int base = 0; base < B * CP; base += 512) {
      float v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int i = base + k * 128 + et;
        const int row = i / CP, c = i - row * CP;
        v[k] = (i < B * CP && c < C) ? __ldcg(a.logits + (size_t)row * a.ld_logits + c) : -INFINITY;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int i = base + k * 128 + et;
        if (i < B * CP) xs[i] = v[k];
      }
    }
    if (et < B) lab[et] = lab_mine;
    asm volatile("bar.sync 1, 128;" ::: "memory");
    CE_STAMP(1);
    float loss_row = 0.f;
    if (et < B) {
      const float4* xr4 = reinterpret_cast<const float4*>(xs + et * CP);
      float mx = -INFINITY;
      for (int c4 = 0; c4 < CP / 4; ++c4) {
        const float4 q4 = xr4[c4];
        mx = fmaxf(fmaxf(mx, fmaxf(q4.x, q4.y)), fmaxf(q4.z, q4.w));
      }
      float se = 0.f;
      for (int c4 = 0; c4 < CP / 4; ++c4) {
        const float4 q4 = xr4[c4];
        se += (__expf(q4.x - mx) + __expf(q4.y - mx)) + (__expf(q4.z - mx) + __expf(q4.w - mx));  // exp(-inf) = 0
      }
      const float lse = mx + __logf(se);
      s_lse[et] = lse;
      if (lab_mine >= 0 && lab_mine < C) loss_row = lse - xs[et * CP + lab_mine];
    }
    loss_row = warp_sum(loss_row);
    if (a.loss_sum && lane == 0 && loss_row != 0.f) atomicAdd(a.loss_sum, loss_row);
    asm volatile("bar.sync 1, 128;" ::: "memory");
    CE_STAMP(2);
    for (int i = et; i < B * C; i += 128) {
      const int row = i / C, c = i - row * C;
      const __nv_bfloat16 gb =
          __float2bfloat16_rn((__expf(xs[row * CP + c] - s_lse[row]) - (c == lab[row] ? 1.f : 0.f)) * inv_b);
      if (dl) dl[(size_t)row * a.ld_d + c] = gb;
      if (dlt) dlt[(size_t)c * a.ld_dt + row] = gb;
      gs[i] = __bfloat162float(gb);
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    CE_STAMP(3);
    if (a.dbias)
      for (int c = et; c < C; c += 128) {
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        int r2 = 0;
        for (; r2 + 4 <= B; r2 += 4) {
          s0 += gs[(r2 + 0) * C + c];
          s1 += gs[(r2 + 1) * C + c];
          s2 += gs[(r2 + 2) * C + c];
          s3 += gs[(r2 + 3) * C + c];
        }
        for (; r2 < B; ++r2) s0 += gs[r2 * C + c];
        a.dbias[c] = (s0 + s1) + (s2 + s3);
      }
    CE_STAMP(4);
    return;
  }
#undef CE_STAMP
  for (int row = ewarp; row < B; row += 4) {
    // the logits may have been stored by another CTA of the cluster: never through L1
    const float* xg = a.logits + (size_t)row * a.ld_logits;
    const float* xs = scratch + (size_t)row * C;
    auto X = [&](int c) { return scratch ? xs[c] : __ldcg(xg + c); };
    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) mx = fmaxf(mx, X(c));
    mx = warp_max(mx);
    float se = 0.f;
    for (int c = lane; c < C; c += 32) se += __expf(X(c) - mx);
    se = warp_sum(se);
    const float lse = mx + __logf(se);
    const int y = scratch ? s_lab[row] : a.labels[row];
    for (int c = lane; c < C; c += 32) {
      const float g = (__expf(X(c) - lse) - (c == y ? 1.f : 0.f)) * inv_b;
      const __nv_bfloat16 gb = __float2bfloat16_rn(g);
      if (dl) dl[(size_t)row * a.ld_d + c] = gb;
      if (dlt) dlt[(size_t)c * a.ld_dt + row] = gb;
      atomicAdd(s_db + c, __bfloat162float(gb));
    }
    if (lane == 0 && y >= 0 && y < C) wloss += lse - X(y);
  }
  asm volatile("bar.sync 1, 128;" ::: "memory");
  if (a.dbias)
    for (int c = et; c < C; c += 128) a.dbias[c] = s_db[c];
  if (a.loss_sum && lane == 0 && wloss != 0.f) atomicAdd(a.loss_sum, wloss);
}

__global__ void __launch_bounds__(kThreadsGemm, 1)
gemm_chain_kernel(const __grid_constant__ ChainArgs ga) {
  __shared__ ChainScalars g;  // see ChainScalars: one parallel copy instead of serial constant misses
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&ga.sc);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&g);
    for (int i = threadIdx.x; i < (int)(sizeof(ChainScalars) / 4); i += kThreadsGemm) dst[i] = src[i];
  }
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t full_bar[kChainStages];
  __shared__ __align__(8) uint64_t empty_bar[kChainStages];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_smem;
  __shared__ float s_db[256];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < kChainMaxGemm; ++i)
      if (ga.sc.M[i] > 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&ga.ma[i])) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&ga.mb[i])) : "memory");
      }
    for (int s = 0; s < kChainStages; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&tmem_full_bar[a]), 1);
      mbar_init(smem_u32(&tmem_empty_bar[a]), 4);  // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_smem), 2 * kChainMaxBN);
  pdl_launch_dependents();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();

  uint32_t it = 0;       // k-block counter over the whole chain (smem ring position)
  uint32_t tile_it = 0;  // output-tile counter over the whole chain (TMEM accumulator position)
  // launched as ONE thread-block cluster: the tiles of a link are dealt round-robin to the CTAs,
  // links are separated by the hardware cluster barrier (release / acquire)
  const int crank = (int)cluster_ctarank(), csize = (int)cluster_nctarank();

#ifdef P2PFL_BRINGUP
  long long* const stamps = (g.stamps && crank == 0) ? g.stamps : nullptr;
#define CHAIN_STAMP(slot) \
  if (stamps && lane == 0) stamps[p * 8 + (slot)] = clock64()
#else
#define CHAIN_STAMP(slot)
#endif
  for (int p = 0; p < g.n_phases; ++p) {
    if (threadIdx.x == 0) CHAIN_STAMP(0);
    if (g.kind[p] == 0) {
      const int gi0 = g.idx[p], cnt = g.cnt[p] > 0 ? g.cnt[p] : 1;
      int n_tiles = 0;
      for (int j = 0; j < cnt; ++j) n_tiles += chain_tiles_of(g, gi0 + j);
      // what this CTA will do in this link (every role advances the ring / accumulator counters alike)
      uint32_t my_tiles = 0, my_kb = 0;
      for (int t = crank; t < n_tiles; t += csize) {
        ++my_tiles;
        my_kb += (uint32_t)chain_locate(g, gi0, cnt, t).num_kb;
      }
      if (warp == 0) {
        // ---------------- TMA producer ----------------
        if (elect_one()) {
          uint32_t lit = it;
          for (int t = crank; t < n_tiles; t += csize) {
            const ChainTile ct = chain_locate(g, gi0, cnt, t);
            for (int kb = 0; kb < ct.num_kb; ++kb, ++lit) {
              const uint32_t s = lit % kChainStages, ph = (lit / kChainStages) & 1u;
              mbar_wait(smem_u32(&empty_bar[s]), ph ^ 1u);
              const uint32_t bar = smem_u32(&full_bar[s]);
              mbar_expect_tx(bar, (uint32_t)(ct.a_rows + ct.bn) * BK * 2);
              const uint32_t dst = smem_u32(smem + (size_t)s * kChainStageBytes);
              tma_load_2d(&ga.ma[ct.gi], bar, dst, kb * BK, ct.m0);
              tma_load_2d(&ga.mb[ct.gi], bar, dst + kTileABytes, kb * BK, ct.n0);
            }
          }
        }
        __syncwarp();
      } else if (warp == 1) {
        // ---------------- MMA issuer ----------------
        uint32_t lit = it, ltile = tile_it;
        for (int t = crank; t < n_tiles; t += csize, ++ltile) {
          const ChainTile ct = chain_locate(g, gi0, cnt, t);
          const uint32_t idesc = make_idesc((uint32_t)ct.bn);
          const uint32_t acc = ltile & 1u, aph = (ltile >> 1) & 1u;
          mbar_wait(smem_u32(&tmem_empty_bar[acc]), aph ^ 1u);  // epilogue drained this accumulator
          tc_fence_after();
          for (int kb = 0; kb < ct.num_kb; ++kb, ++lit) {
            const uint32_t s = lit % kChainStages, ph = (lit / kChainStages) & 1u;
            mbar_wait(smem_u32(&full_bar[s]), ph);
            tc_fence_after();
            if (kb == 0 && t == crank) CHAIN_STAMP(1);  // first operands landed
            if (elect_one()) {
              const uint32_t a_addr = smem_u32(smem + (size_t)s * kChainStageBytes);
              const uint64_t da = make_smem_desc(a_addr);
              const uint64_t db = make_smem_desc(a_addr + kTileABytes);
#pragma unroll
              for (int k = 0; k < BK / 16; ++k)
                umma_bf16(tmem_base + acc * kChainMaxBN, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc,
                          (kb | k) != 0 ? 1u : 0u);
              umma_commit(smem_u32(&empty_bar[s]));
              if (kb == ct.num_kb - 1) umma_commit(smem_u32(&tmem_full_bar[acc]));
            }
            __syncwarp();
          }
          CHAIN_STAMP(2);  // this tile's MMAs issued
        }
      } else {
        // ---------------- epilogue ----------------
        const int q = warp & 3;
        uint32_t ltile = tile_it;
        for (int t = crank; t < n_tiles; t += csize, ++ltile) {
          const ChainTile ct = chain_locate(g, gi0, cnt, t);
          const uint32_t acc = ltile & 1u, aph = (ltile >> 1) & 1u;
          mbar_wait(smem_u32(&tmem_full_bar[acc]), aph);
          tc_fence_after();
          if (warp == 2 && t == crank) CHAIN_STAMP(3);  // first accumulator complete
          epilogue_tile(g.ep[ct.gi], g.M[ct.gi], g.N[ct.gi], ct.m0, ct.n0, (uint32_t)ct.bn, tmem_base + acc * kChainMaxBN, q,
                        lane, g.N[ct.gi] <= ct.bn
#ifdef P2PFL_BRINGUP
                        , (stamps && warp == 2 && t == crank) ? stamps + 96 + p * 8 : nullptr
#endif
          );
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&tmem_empty_bar[acc]));
        }
        if (warp == 2) CHAIN_STAMP(4);  // stores issued
      }
      tile_it += my_tiles;
      it += my_kb;
    } else if (warp >= 2 && crank == 0) {
      const bool fits = (size_t)g.ce.B * (2 * g.ce.C + 6) * 4 <= (size_t)kChainStages * kChainStageBytes;
      chain_ce(g.ce, warp - 2, lane, s_db, fits ? reinterpret_cast<float*>(smem) : nullptr, g.ce.B <= 128 && g.ce.C <= 64
#ifdef P2PFL_BRINGUP
               , stamps ? stamps + 96 + p * 8 : nullptr
#endif
      );
    }
    // ---- link boundary: this link's global stores (generic proxy, possibly another CTA's) must
    // be visible to the next link's TMA loads (async proxy) and to its plain loads
    // (the cluster barrier's release / acquire pair orders the stores at cluster scope itself)
    if (csize == 1) __threadfence();
    if (warp >= 2) asm volatile("fence.proxy.async.global;" ::: "memory");  // the threads that stored
    if (warp == 2) CHAIN_STAMP(5);  // fences done
    if (csize > 1) cluster_sync_release_acquire();
    else __syncthreads();
    if (warp == 0) asm volatile("fence.proxy.async.global;" ::: "memory");  // the thread that issues TMA loads
    if (warp == 2) CHAIN_STAMP(6);  // barrier passed
  }
#undef CHAIN_STAMP

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * kChainMaxBN);
  }
}

}  // namespace

// Largest cluster (16, 8, 4, 2, 1) this device can co-schedule for the chain kernel.
extern "C" int p2pfl_gemm_chain_max_cluster(int device) {
  if (cudaSetDevice(device) != cudaSuccess) return 1;
  constexpr size_t smem = (size_t)kChainStages * kChainStageBytes + 1024;
  cudaFuncSetAttribute(gemm_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(gemm_chain_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  for (int c = 16; c > 1; c >>= 1) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(c);
    cfg.blockDim = dim3(kThreadsGemm);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)c;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, gemm_chain_kernel, &cfg) == cudaSuccess && n >= 1) return c;
    (void)cudaGetLastError();
  }
  return 1;
}

#ifdef P2PFL_BRINGUP
static long long* g_chain_stamps = nullptr;
#endif
// Per-link clock stamps exist only in builds with -DP2PFL_BRINGUP; the shipped library refuses.
extern "C" int p2pfl_gemm_chain_set_stamps(void* device_buffer) {
#ifdef P2PFL_BRINGUP
  g_chain_stamps = reinterpret_cast<long long*>(device_buffer);
  return 0;
#else
  if (device_buffer == nullptr) return 0;
  set_error("clock stamps are compiled out of this build (rebuild with -DP2PFL_BRINGUP)");
  return -1;
#endif
}

extern "C" int p2pfl_gemm_chain(int device, void* stream, const p2pfl_gemm_args* gemms, int n_gemm,
                                const int32_t* phase_kind, const int32_t* phase_idx, int n_phases,
                                const p2pfl_ce_args* ce) {
  return p2pfl_gemm_chain_cluster(device, stream, gemms, n_gemm, phase_kind, phase_idx, nullptr, n_phases, ce, 1);
}

extern "C" int p2pfl_gemm_chain_cluster(int device, void* stream, const p2pfl_gemm_args* gemms, int n_gemm,
                                        const int32_t* phase_kind, const int32_t* phase_idx, const int32_t* phase_cnt,
                                        int n_phases, const p2pfl_ce_args* ce, int cluster_ctas) {
  P2PFL_REQUIRE(cluster_ctas >= 1 && cluster_ctas <= 16 && (cluster_ctas & (cluster_ctas - 1)) == 0,
                "the chain runs as one cluster of 1, 2, 4, 8 or 16 CTAs");
  P2PFL_REQUIRE(n_gemm >= 0 && n_gemm <= kChainMaxGemm && n_phases >= 1 && n_phases <= kChainMaxPhases,
                "a GEMM chain takes at most 8 GEMMs in at most 12 links");
  P2PFL_REQUIRE(phase_kind && phase_idx, "null phase list");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  ChainArgs g;
  memset(&g, 0, sizeof(g));
  GroupedArgs tmp;  // reuse the single-problem filler (maps + epilogue), then copy across
  for (int i = 0; i < n_gemm; ++i) {
    const p2pfl_gemm_args& a = gemms[i];
    P2PFL_REQUIRE(a.A && a.B && a.M > 0 && a.N > 0 && a.K > 0, "bad GEMM shape or operands");
    P2PFL_REQUIRE((a.split_k == 0 || a.split_k == 1) && a.mn_major == 0, "chain links are plain K-major GEMMs without split-K");
    memset(&tmp, 0, sizeof(tmp));
    if (int rc = fill_problem(tmp, 0, a)) return rc;
    // a batch-major link has M = batch rows: fetch only those (the MMA still multiplies 128 rows, the
    // rest of the tile holds whatever was there and lands in accumulator rows nobody stores)
    const int a_rows = a.M >= BM ? BM : (a.M + 7) / 8 * 8;
    g.sc.a_rows[i] = a_rows;
    if (int rc = make_map(&g.ma[i], a.A, a.M, a.K, a.lda, a_rows)) return rc;
    if (int rc = make_map(&g.mb[i], a.B, a.N, a.K, a.ldb, chain_bn_for(a.N))) return rc;  // the chain's own N tiling
    g.sc.ep[i] = tmp.ep[0];
    g.sc.M[i] = a.M;
    g.sc.N[i] = a.N;
    g.sc.K[i] = a.K;
    const int cbn = chain_bn_for(a.N);
    g.sc.tiles_n[i] = (a.N + cbn - 1) / cbn;
    g.sc.tiles[i] = ((a.M + BM - 1) / BM) * g.sc.tiles_n[i];
    g.sc.num_kb[i] = (a.K + BK - 1) / BK;
  }
  bool has_ce = false;
  for (int p = 0; p < n_phases; ++p) {
    P2PFL_REQUIRE(phase_kind[p] == 0 || phase_kind[p] == 1, "unknown link kind");
    const int cnt = (phase_cnt && phase_kind[p] == 0 && phase_cnt[p] > 0) ? phase_cnt[p] : 1;
    P2PFL_REQUIRE(phase_kind[p] == 1 || (phase_idx[p] >= 0 && phase_idx[p] + cnt <= n_gemm), "link refers to a missing GEMM");
    g.sc.kind[p] = phase_kind[p];
    g.sc.idx[p] = phase_idx[p];
    g.sc.cnt[p] = cnt;
    has_ce |= phase_kind[p] == 1;
  }
  g.sc.n_phases = n_phases;
#ifdef P2PFL_BRINGUP
  g.sc.stamps = g_chain_stamps;
#endif
  if (has_ce) {
    P2PFL_REQUIRE(ce && ce->logits && ce->labels && ce->B > 0 && ce->C > 0 && ce->C <= 256, "bad cross-entropy link");
    g.sc.ce = *ce;
  }
  constexpr size_t smem = (size_t)kChainStages * kChainStageBytes + 1024;
  static bool attr_set[64] = {false};
  if (device >= 0 && device < 64 && !attr_set[device]) {
    P2PFL_CUDA_OK(cudaFuncSetAttribute(gemm_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set[device] = true;
  }
  if (cluster_ctas > 8) {
    static bool np_set[64] = {false};
    if (device >= 0 && device < 64 && !np_set[device]) {
      P2PFL_CUDA_OK(cudaFuncSetAttribute(gemm_chain_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
      np_set[device] = true;
    }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cluster_ctas);
  cfg.blockDim = dim3(kThreadsGemm);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = as_stream(stream);
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)cluster_ctas;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  P2PFL_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_chain_kernel, g));
  return 0;
}
