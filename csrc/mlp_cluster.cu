// A whole training EPOCH of the MNIST-shaped MLP (784-256-128-10, batch <= 32) as ONE launch of
// ONE thread-block cluster of 16 CTAs.
//
// reference sites replaced: mlp.py:57-80 (forward, loss, Adam) + autograd's backward, i.e. what
// Lightning's Trainer.fit does per batch (lightning_learner.py:180-214), for every batch of the epoch.
//
// Why: the step is ~45 MFLOP - all latency.  As a chain of GEMM launches (or as the cluster chain of
// gemm_tcgen05.cu) every link pays a TMA round trip to L2 for its operands, a global store of its
// result, a proxy fence and a cluster barrier: ~4.4 us per link, 37 us per step.  Here nothing but
// the batch itself is fetched per step:
//   * the bf16 weights are RESIDENT in shared memory for the whole epoch: CTA c keeps rows
//     16c..16c+15 of W1 (26 KB) and a full copy of W2 and W3 (68 KB);
//   * layer 1 is split over the 16 CTAs (16 output features each); the 32 x 16 slices of H1 go to
//     every CTA through distributed shared memory, already in the tcgen05 operand layout;
//   * layers 2 and 3, the loss and the two data-gradient GEMMs are computed by EVERY CTA (they are
//     tiny), so the forward/backward chain needs one cluster barrier, not one per link, and each CTA
//     ends up with the slice of dZ1 that its rows of W1 need;
//   * weight gradients come out of TMEM in the CTA that owns the rows and are consumed there by
//     Adam (fp32 master / moments in global memory, L2 resident); the owner rewrites its bf16 rows
//     in shared memory (W1) or in global memory (W2, W3: every CTA re-reads them by TMA, which
//     overlaps the next step's layer 1).
// All GEMMs are tcgen05.mma (kind::f16, bf16 x bf16 -> fp32 in TMEM), M = 128 with the batch (or the
// input-feature index, for weight gradients) on the accumulator lanes; operands are 128B-swizzled
// boxes of [rows x 64 columns], read K-major or MN-major as the link needs - one copy serves both.
#include <string.h>

#include "common.cuh"
#include "tcgen05_ptx.cuh"

using namespace p2pfl;
using namespace p2pfl::tc;

namespace {

constexpr int kC = 16;  // CTAs in the cluster
constexpr int D0 = 784, D1 = 256, D2 = 128, D3 = 10;
constexpr int kRows = 32;            // batch rows per step (boxes are sized for it)
constexpr int kK0Boxes = 13;         // 64-wide column boxes of the layer-1 input (784 -> 832, zero padded)
constexpr int kR1 = D1 / kC;         // rows of W1 (= features of layer 1) per CTA
constexpr int kR2 = D2 / kC;         // rows of W2 per CTA (8; their GEMM is 16 wide, the upper half is ignored)
constexpr int kW3Owner = 8;          // CTA 8 owns W3 / b3
constexpr int kWarps = 16;           // four warps per scheduler: 128 registers per thread
constexpr int kIssuer = 1;           // lane 0 of this warp issues every TMA load and every MMA
constexpr int kThreads = kWarps * 32;

// offsets in the flat parameter vector (ParamLayout: state_dict order, 128-byte aligned slots)
constexpr uint32_t kW1Off = 0, kB1Off = kW1Off + D1 * D0, kW2Off = kB1Off + D1, kB2Off = kW2Off + D2 * D1,
                   kW3Off = kB2Off + D2, kB3Off = kW3Off + D3 * D2, kFlatTotal = kB3Off + 32;
static_assert(kB1Off % 32 == 0 && kW2Off % 32 == 0 && kB2Off % 32 == 0 && kW3Off % 32 == 0 && kB3Off % 32 == 0,
              "slots start on 128-byte boundaries");

// shared-memory map (bytes from a 1024-aligned base); every box is [rows x 128 B], 128B swizzle.
// Activation boxes hold 32 rows but are read as M = 128 operands: rows 32..127 alias whatever
// follows (their accumulator lanes are never read), so the order below keeps those reads in bounds.
constexpr uint32_t kBoxX = kRows * 128;                   // 4 KB
constexpr uint32_t oXB = 0;                               // X      [32 x 832]  13 boxes
constexpr uint32_t oH1 = oXB + kK0Boxes * kBoxX;          // H1     [32 x 256]   4 boxes
constexpr uint32_t oH2 = oH1 + 4 * kBoxX;                 // H2     [32 x 128]   2 boxes
constexpr uint32_t oDZ3 = oH2 + 2 * kBoxX;                // dZ3    [32 x 16(64)]
constexpr uint32_t oDZ2 = oDZ3 + kBoxX;                   // dZ2    [32 x 128]   2 boxes
constexpr uint32_t oDZ1S = oDZ2 + 2 * kBoxX;              // dZ1, own 16 columns
constexpr uint32_t kBoxW1 = kR1 * 128;                    // 2 KB
constexpr uint32_t oW1 = oDZ1S + kBoxX;                   // W1 rows of this CTA [16 x 832] 13 boxes
constexpr uint32_t kBoxW2 = D2 * 128;                     // 16 KB
constexpr uint32_t oW2 = oW1 + kK0Boxes * kBoxW1;         // W2 [128 x 256] 4 boxes
constexpr uint32_t kBoxW3 = 16 * 128;                     // 2 KB
constexpr uint32_t oW3 = oW2 + 4 * kBoxW2;                // W3 [16 x 128] 2 boxes (rows 10..15 zero)
constexpr uint32_t kSmemBytes = oW3 + 2 * kBoxW3;
static_assert(kSmemBytes <= 200 * 1024, "shared-memory budget");
static_assert(oW1 % 1024 == 0 && oW2 % 1024 == 0 && oW3 % 1024 == 0, "swizzled boxes start on 1 KB boundaries");

// TMEM columns.  Accumulators whose lifetimes do not overlap share columns; the 336 columns that
// are left hold the fp32 master and Adam's two moments of this CTA's W1 rows for the whole epoch, in
// the lane / column map of the weight-gradient accumulator (lane = input feature mod 128, column =
// 16 * tile + row): the update of W1 is TMEM -> registers -> TMEM, with no other memory traffic.
constexpr uint32_t cFwd2 = 0, cDg3 = 0, cWg2 = 0;                 // [128 features x 32 batch columns] / 2 tiles x 16
constexpr uint32_t cFwd1 = 32, cFwd3 = 32, cDg2 = 32, cWg3 = 32;  // [32 batch rows x 16] / 1 tile x 16
constexpr uint32_t cWg1 = 48;                                     // 7 tiles x 16 columns
constexpr uint32_t cM = 160, cV = 272, cP = 384;                  // 7 x 16 columns each
static_assert(cP + 112 <= 512, "TMEM columns");

struct Args {
  const int32_t* ys;
  float* flat;
  float* m;
  float* v;
  __nv_bfloat16* w16;
  float* grads;  // nullable: every gradient is also written there (tests)
  float* loss;   // [n_steps], += sum of the per-sample losses of the step
  const int32_t* step_ptr;
  int32_t n_rows, n_steps;
  float lr, b1, b2, eps;
#ifdef P2PFL_BRINGUP
  long long* stamps;  // [16 CTAs][32] SM-clock stamps of one steady-state step, see p2pfl_mlp_epoch_set_stamps
#endif
};

// byte offset of element (row, col) inside a swizzled box row group: col in [0, 64)
__device__ __forceinline__ uint32_t swz(uint32_t row, uint32_t col) {
  return row * 128u + ((((col >> 3) ^ row) & 7u) << 4) + ((col & 7u) << 1);
}

// MN-major descriptor over [k rows x 64-column boxes]: LBO = distance between column boxes
__device__ __forceinline__ uint64_t mn_desc(uint32_t addr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(lbo_bytes >> 4) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ float ld_shared_bf16(uint32_t addr) {
  unsigned short h;
  asm volatile("ld.shared.b16 %0, [%1];" : "=h"(h) : "r"(addr));
  return __uint_as_float((uint32_t)h << 16);
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of `local` (a shared::cta address of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_u4(uint32_t addr, const uint4& v) {
  asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void st_shared_u4(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_shared_u4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void st_shared_b16(uint32_t addr, __nv_bfloat16 v) {
  asm volatile("st.shared.b16 [%0], %1;" ::"r"(addr), "h"(*reinterpret_cast<unsigned short*>(&v)) : "memory");
}
// generic-proxy writes to shared memory become visible to the async proxy (tcgen05.mma operands)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_async_all() { asm volatile("fence.proxy.async.shared::cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }

struct AdamK {
  float b1, b2, one_m_b1, one_m_b2, step_size, inv_sqrt_bc2, eps;
};
__device__ __forceinline__ AdamK adam_consts(const Args& a, int t) {
  AdamK k;
  k.b1 = a.b1;
  k.b2 = a.b2;
  k.one_m_b1 = 1.f - a.b1;
  k.one_m_b2 = 1.f - a.b2;
  const float bc1 = 1.f - powf(a.b1, (float)t), bc2 = 1.f - powf(a.b2, (float)t);
  k.step_size = a.lr / bc1;
  k.inv_sqrt_bc2 = rsqrtf(bc2);
  k.eps = a.eps;
  return k;
}
// Same formula as adam_kernel (elementwise.cu), weight decay 0, with the hardware's approximate
// square root and reciprocal (2 ulp each): the 16 SMs of the cluster update all 235 k parameters every
// step, and with IEEE sqrt / division this loop alone was 7 us of a 21 us step.
__device__ __forceinline__ float adam_one(const AdamK& k, float g, float& p, float& m, float& v) {
  m = fmaf(k.b1, m, k.one_m_b1 * g);
  v = fmaf(k.b2, v, k.one_m_b2 * g * g);
  float sq, rc;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(v));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(fmaf(sq, k.inv_sqrt_bc2, k.eps)));
  p = fmaf(-k.step_size * m, rc, p);
  return p;
}

// 32 fp32 accumulator columns of the thread's row -> bias -> (ReLU) -> 32 bf16 = four 16-byte chunks
__device__ __forceinline__ void pack32(const float (&x)[32], uint4 (&out)[4]) {
#pragma unroll
  for (int ch = 0; ch < 4; ++ch) {
    out[ch].x = pack_bf16x2(x[ch * 8 + 0], x[ch * 8 + 1]);
    out[ch].y = pack_bf16x2(x[ch * 8 + 2], x[ch * 8 + 3]);
    out[ch].z = pack_bf16x2(x[ch * 8 + 4], x[ch * 8 + 5]);
    out[ch].w = pack_bf16x2(x[ch * 8 + 6], x[ch * 8 + 7]);
  }
}
__device__ __forceinline__ void pack16(const float (&x)[16], uint4 (&out)[2]) {
#pragma unroll
  for (int ch = 0; ch < 2; ++ch) {
    out[ch].x = pack_bf16x2(x[ch * 8 + 0], x[ch * 8 + 1]);
    out[ch].y = pack_bf16x2(x[ch * 8 + 2], x[ch * 8 + 3]);
    out[ch].z = pack_bf16x2(x[ch * 8 + 4], x[ch * 8 + 5]);
    out[ch].w = pack_bf16x2(x[ch * 8 + 6], x[ch * 8 + 7]);
  }
}
// address of 16-byte chunk `chunk` (0..7) of row `row` in a swizzled box
__device__ __forceinline__ uint32_t chunk_addr(uint32_t box, uint32_t row, uint32_t chunk) {
  return box + row * 128u + (((chunk ^ row) & 7u) << 4);
}

// Adam over 16 elements that one thread owns: element j lives at flat index base + j * stride
// (stride = row length: consecutive lanes hold consecutive columns, so every access coalesces).
template <int kCount>
__device__ __forceinline__ void adam_column(const Args& a, const AdamK& k, const uint32_t (&r)[16], uint32_t base, uint32_t stride,
                                            float (&p_out)[16]) {
  float p[16], m[16], v[16];
#pragma unroll
  for (int j = 0; j < kCount; ++j) {
    const uint32_t idx = base + (uint32_t)j * stride;
    p[j] = __ldcg(a.flat + idx);
    m[j] = __ldcg(a.m + idx);
    v[j] = __ldcg(a.v + idx);
  }
#pragma unroll
  for (int j = 0; j < kCount; ++j) {
    const uint32_t idx = base + (uint32_t)j * stride;
    const float g = __uint_as_float(r[j]);
    p_out[j] = adam_one(k, g, p[j], m[j], v[j]);
    a.flat[idx] = p[j];
    a.m[idx] = m[j];
    a.v[idx] = v[j];
    a.w16[idx] = __float2bfloat16_rn(p[j]);
    if (a.grads) a.grads[idx] = g;
  }
}
// Adam of one bias element (gradient already summed over the batch)
__device__ __forceinline__ float adam_bias(const Args& a, const AdamK& k, uint32_t idx, float g) {
  float p = __ldcg(a.flat + idx), m = __ldcg(a.m + idx), v = __ldcg(a.v + idx);
  adam_one(k, g, p, m, v);
  a.flat[idx] = p;
  a.m[idx] = m;
  a.v[idx] = v;
  a.w16[idx] = __float2bfloat16_rn(p);
  if (a.grads) a.grads[idx] = g;
  return p;
}

__global__ void __launch_bounds__(kThreads, 1)
mlp_epoch_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW1,
                 const __grid_constant__ CUtensorMap tmW2, const __grid_constant__ CUtensorMap tmW3, const Args a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t sb = smem_u32(smem);
  __shared__ __align__(8) uint64_t mb_mma, mb_x, mb_w, mb_w1;
  __shared__ uint32_t tmem_base_smem;
  __shared__ float s_b1[kR1];      // this CTA's slice of b1 (kept current by the lanes that update it)
  __shared__ float s_db2[4][kR2];  // batch sums of this CTA's 8 columns of dZ2: one partial per warp group

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t c = cluster_rank();
  const bool issuer = warp == kIssuer;
  const bool one = issuer && lane == 0;  // the thread that issues every TMA load and every MMA
  const int q = warp & 3, g = warp >> 2;  // TMEM lane quarter of this warp / its group of four
  const bool q0 = q == 0;                 // warps that can read accumulator lanes 0..31 (the batch rows)
  const uint32_t lane_base = (uint32_t)(q * 32) << 16;

  if (one) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmX)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmW2)) : "memory");
    mbar_init(smem_u32(&mb_mma), 1);
    mbar_init(smem_u32(&mb_x), 1);
    mbar_init(smem_u32(&mb_w), 1);
    mbar_init(smem_u32(&mb_w1), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_smem), 512);
  if (threadIdx.x >= 32 && threadIdx.x < 32 + kR1) s_b1[threadIdx.x - 32] = a.flat[kB1Off + kR1 * c + threadIdx.x - 32];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_smem;

  if (one) {
    // resident operands: this CTA's rows of W1, and W2 / W3 whole; then the first batch
    mbar_expect_tx(smem_u32(&mb_w1), kK0Boxes * kBoxW1);
    for (int j = 0; j < kK0Boxes; ++j) tma_load_2d(&tmW1, smem_u32(&mb_w1), sb + oW1 + j * kBoxW1, 64 * j, (int)(kR1 * c));
    mbar_expect_tx(smem_u32(&mb_w), 4 * kBoxW2 + 2 * kBoxW3);
    for (int j = 0; j < 4; ++j) tma_load_2d(&tmW2, smem_u32(&mb_w), sb + oW2 + j * kBoxW2, 64 * j, 0);
    for (int j = 0; j < 2; ++j) tma_load_2d(&tmW3, smem_u32(&mb_w), sb + oW3 + j * kBoxW3, 64 * j, 0);
    mbar_expect_tx(smem_u32(&mb_x), kK0Boxes * kBoxX);
    for (int j = 0; j < kK0Boxes; ++j) tma_load_2d(&tmX, smem_u32(&mb_x), sb + oXB + j * kBoxX, 64 * j, 0);
  }
  // The fp32 master and Adam's moments of this CTA's W1 rows move into TMEM for the whole epoch: lane
  // = input feature (mod 128), column = 16 * tile + row, exactly where the gradient of that weight lands
  for (int tt = g; tt < 7; tt += 4) {
    const uint32_t i = 128u * tt + 32u * q + lane;
    uint32_t rm[16], rv[16], rp[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const uint32_t idx = kW1Off + (kR1 * c + j) * D0 + i;
      rm[j] = i < (uint32_t)D0 ? __float_as_uint(a.m[idx]) : 0u;
      rv[j] = i < (uint32_t)D0 ? __float_as_uint(a.v[idx]) : 0u;
      rp[j] = i < (uint32_t)D0 ? __float_as_uint(a.flat[idx]) : 0u;
    }
    tmem_st16(tmem + lane_base + cM + 16 * tt, rm);
    tmem_st16(tmem + lane_base + cV + 16 * tt, rv);
    tmem_st16(tmem + lane_base + cP + 16 * tt, rp);
  }
  tmem_st_wait();
  tc_fence_before();
  cluster_barrier();  // every CTA of the cluster is running before anybody stores into a neighbour
  tc_fence_after();

  // Layer 1 of step `step` (issuing thread only): H1[:, 16c .. 16c+15] pre-activation = X . W1_c^T.  It
  // needs nothing from the other CTAs - the batch and this CTA's own rows of W1 - so it is issued as
  // soon as Adam has rewritten those rows, BEFORE the step's closing cluster barrier, and runs under it.
  auto issue_layer1 = [&](int step) {
    if (step == 0) mbar_wait(smem_u32(&mb_w1), 0);
    mbar_wait(smem_u32(&mb_x), (uint32_t)step & 1u);
    tc_fence_after();
    fence_async_smem();
    const uint32_t idesc = make_idesc(16);
    for (int j = 0; j < kK0Boxes; ++j) {
      const uint64_t da = make_smem_desc(sb + oXB + j * kBoxX), db = make_smem_desc(sb + oW1 + j * kBoxW1);
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) umma_bf16(tmem + cFwd1, da + 2 * ks, db + 2 * ks, idesc, (j | ks) != 0 ? 1u : 0u);
    }
    umma_commit(smem_u32(&mb_mma));
  };
  if (one) issue_layer1(0);

  const int32_t step0 = a.step_ptr ? *a.step_ptr : 0;
  uint32_t par = 0;  // parity of the next completion of mb_mma (every thread counts the links alike)
  uint32_t hbits = 0;  // warp 0: which of this row's 16 layer-1 activations are positive (ReLU mask)
  // bias gradients (batch sums) wait in a register of the lane that will apply them with the weights
  float g_b1 = 0.f, g_b3 = 0.f;

#define LINK_WAIT()                  \
  mbar_wait(smem_u32(&mb_mma), par); \
  tc_fence_after()
#define LINK_END()    \
  par ^= 1u;          \
  tc_fence_before();  \
  __syncthreads();    \
  tc_fence_after()

#ifdef P2PFL_BRINGUP
#define STAMP(slot) \
  if (a.stamps && warp == 0 && lane == 0 && t == stamp_step) a.stamps[c * 32 + (slot)] = clock64()
  const int stamp_step = a.n_steps > 8 ? a.n_steps / 2 : a.n_steps - 1;
#define STAMP1(slot) \
  if (a.stamps && lane == 0 && t == stamp_step) a.stamps[c * 32 + (slot)] = clock64()
#else
#define STAMP(slot)
#define STAMP1(slot)
#endif
  for (int t = 0; t < a.n_steps; ++t) {
    STAMP(0);
    const int row0 = t * kRows;
    const int bsz = (a.n_rows - row0) < kRows ? (a.n_rows - row0) : kRows;
    const AdamK ak = adam_consts(a, step0 + t + 1);
    const float bias2 = __ldcg(a.flat + kB2Off + 32u * q + lane);  // used after barrier (a): latency hidden
    float bias3[D3];
    if (warp == 0) {
#pragma unroll
      for (int j = 0; j < D3; ++j) bias3[j] = __ldcg(a.flat + kB3Off + j);
    }

    // ================= layer 1: H1[:, 16c .. 16c+15] = relu(X . W1_c^T + b1_c) =================
    // (the MMAs were issued before the previous step's closing barrier - see issue_layer1 below)
    if (q0) {
      float bias[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) bias[j] = s_b1[j];
      LINK_WAIT();
      STAMP(1);
      uint32_t r[16];
      tmem_ld16(tmem + cFwd1, r);
      float h[16];
      uint32_t bits = 0;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        h[j] = fmaxf(__uint_as_float(r[j]) + bias[j], 0.f);
        bits |= (h[j] > 0.f ? 1u : 0u) << j;
      }
      hbits = bits;
      uint4 u[2];
      pack16(h, u);
      // the slice goes to EVERY CTA's H1 (each of the four row-reading warps serves four CTAs)
      const uint32_t box = sb + oH1 + (c >> 2) * kBoxX;
      const uint32_t a0 = chunk_addr(box, lane, 2 * (c & 3)), a1 = chunk_addr(box, lane, 2 * (c & 3) + 1);
#pragma unroll
      for (int d = 0; d < 4; ++d) {
        const uint32_t dst = (uint32_t)(g * 4 + d);
        st_cluster_u4(map_to_cta(a0, dst), u[0]);
        st_cluster_u4(map_to_cta(a1, dst), u[1]);
      }
      fence_async_all();  // these stores (other CTAs' shared memory included) vs. the MMAs that will read them
      STAMP(2);
    }
    par ^= 1u;
    tc_fence_before();
    cluster_barrier();  // (a) all 16 slices of H1 are in place everywhere
    tc_fence_after();
    STAMP(3);

    // ================= layer 2 (every CTA): H2^T[feature, batch] = relu(W2 . H1^T + b2) =================
    // The 128 features sit on the accumulator lanes, so all four lane quarters hold live rows and
    // all 16 warps share the epilogue: 8 batch columns per thread (with the batch on the lanes only
    // the four quarter-0 warps could read the result, 32 columns per thread, one after the other).
    if (issuer) {
      if (lane == 0) {
        mbar_wait(smem_u32(&mb_w), (uint32_t)t & 1u);  // W2 / W3 as updated by the previous step
        tc_fence_after();
        fence_async_smem();
        const uint32_t idesc = make_idesc(32);
        for (int j = 0; j < 4; ++j) {
          const uint64_t da = make_smem_desc(sb + oW2 + j * kBoxW2), db = make_smem_desc(sb + oH1 + j * kBoxX);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) umma_bf16(tmem + cFwd2, da + 2 * ks, db + 2 * ks, idesc, (j | ks) != 0 ? 1u : 0u);
        }
        umma_commit(smem_u32(&mb_mma));
      }
      __syncwarp();
    }
    {
      const uint32_t o = 32u * q + lane;  // feature of this thread
      const float bias = bias2;
      LINK_WAIT();
      uint32_t r[8];
      tmem_ld8(tmem + lane_base + cFwd2 + 8 * g, r);
      const uint32_t dst = sb + oH2 + (o >> 6) * kBoxX + (((o & 63u) >> 3) << 4) + ((o & 7u) << 1);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t b = 8u * g + k;  // batch row
        st_shared_b16((dst + b * 128u) ^ ((b & 7u) << 4), __float2bfloat16_rn(fmaxf(__uint_as_float(r[k]) + bias, 0.f)));
      }
      fence_async_smem();
    }
    LINK_END();

    STAMP(4);
    // ================= layer 3 + softmax cross-entropy (every CTA): dZ3 =================
    if (issuer) {
      if (lane == 0) {
        const uint32_t idesc = make_idesc(16);
        for (int j = 0; j < 2; ++j) {
          const uint64_t da = make_smem_desc(sb + oH2 + j * kBoxX), db = make_smem_desc(sb + oW3 + j * kBoxW3);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) umma_bf16(tmem + cFwd3, da + 2 * ks, db + 2 * ks, idesc, (j | ks) != 0 ? 1u : 0u);
        }
        umma_commit(smem_u32(&mb_mma));
      }
      __syncwarp();
    } else if (warp == 0) {
      const int y = lane < bsz ? a.ys[row0 + lane] : -1;
      LINK_WAIT();
      uint32_t r[16];
      tmem_ld16(tmem + cFwd3, r);
      float x[D3], mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < D3; ++j) {
        x[j] = __uint_as_float(r[j]) + bias3[j];
        mx = fmaxf(mx, x[j]);
      }
      float se = 0.f;
#pragma unroll
      for (int j = 0; j < D3; ++j) se += __expf(x[j] - mx);
      const float lse = mx + __logf(se);
      const float inv_b = 1.f / (float)bsz;
      float gq[16], loss = 0.f;
#pragma unroll
      for (int j = 0; j < 16; ++j) gq[j] = 0.f;
#pragma unroll
      for (int j = 0; j < D3; ++j) {
        if (lane < bsz) {
          const float gj = (__expf(x[j] - lse) - (j == y ? 1.f : 0.f)) * inv_b;
          gq[j] = __bfloat162float(__float2bfloat16_rn(gj));  // what the backward GEMMs will see
          if (j == y) loss = lse - x[j];
        }
      }
      uint4 u[2];
      pack16(gq, u);
      st_shared_u4(chunk_addr(sb + oDZ3, lane, 0), u[0]);
      st_shared_u4(chunk_addr(sb + oDZ3, lane, 1), u[1]);
      fence_async_smem();
      loss = warp_sum(loss);
      if (c == 0 && lane == 0 && a.loss) a.loss[t] = loss;
      if (c == kW3Owner) {
#pragma unroll
        for (int j = 0; j < D3; ++j) {
          const float s = warp_sum(gq[j]);
          if (lane == j) g_b3 = s;
        }
      }
    }
    LINK_END();

    STAMP(5);
    // ================= dZ2^T[feature, batch] = (W3^T . dZ3^T) * (H2 > 0)  (every CTA) =================
    if (issuer) {
      if (lane == 0) {
        // A = the W3 boxes read MN-major (M = its 128 columns, K = its 16 rows), B = dZ3 (N = batch rows)
        umma_bf16(tmem + cDg3, mn_desc(sb + oW3, kBoxW3), make_smem_desc(sb + oDZ3), make_idesc(32, 1), 0u);
        umma_commit(smem_u32(&mb_mma));
      }
      __syncwarp();
    }
    {
      const uint32_t o = 32u * q + lane;
      const uint32_t off = (o >> 6) * kBoxX + (((o & 63u) >> 3) << 4) + ((o & 7u) << 1);
      float h[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t b = 8u * g + k;
        h[k] = ld_shared_bf16((sb + oH2 + off + b * 128u) ^ ((b & 7u) << 4));
      }
      LINK_WAIT();
      uint32_t r[8];
      tmem_ld8(tmem + lane_base + cDg3 + 8 * g, r);
      float part = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t b = 8u * g + k;
        const float d = h[k] > 0.f ? __uint_as_float(r[k]) : 0.f;
        part += d;
        st_shared_b16((sb + oDZ2 + off + b * 128u) ^ ((b & 7u) << 4), __float2bfloat16_rn(d));
      }
      // this CTA owns rows 8c..8c+7 of W2 (and of b2): their batch sums, one partial per warp group
      if (o - kR2 * c < (uint32_t)kR2) s_db2[g][o - kR2 * c] = part;
      fence_async_smem();
    }
    LINK_END();

    STAMP(6);
    // ================= dZ1[:, 16c .. 16c+15] = (dZ2 . W2)[:, own columns] * (H1 > 0) =================
    if (issuer) {
      if (lane == 0) {
        // K = the 128 output rows of W2, 16 per MMA; B = the W2 boxes read MN-major, only the 16
        // columns this CTA's rows of W1 need (a 32-byte offset inside the swizzled rows of box c / 4)
        const uint32_t idesc = make_idesc(16, 2);
        const uint32_t bsl = sb + oW2 + (c >> 2) * kBoxW2 + (c & 3u) * 32u;
        for (int s = 0; s < 8; ++s) {
          const uint64_t da = make_smem_desc(sb + oDZ2 + (s >> 2) * kBoxX) + 2 * (s & 3);
          const uint64_t db = mn_desc(bsl, kBoxW2) + (uint64_t)s * (2048 >> 4);
          umma_bf16(tmem + cDg2, da, db, idesc, s != 0 ? 1u : 0u);
        }
        umma_commit(smem_u32(&mb_mma));
      }
      __syncwarp();
    } else if (warp == 0) {
      LINK_WAIT();
      uint32_t r[16];
      tmem_ld16(tmem + cDg2, r);
      float x[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) x[j] = (hbits >> j & 1u) ? __uint_as_float(r[j]) : 0.f;
      uint4 u[2];
      pack16(x, u);
      st_shared_u4(chunk_addr(sb + oDZ1S, lane, 0), u[0]);
      st_shared_u4(chunk_addr(sb + oDZ1S, lane, 1), u[1]);
      fence_async_smem();
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float s = warp_sum(x[j]);
        if (lane == j) g_b1 = s;
      }
    }
    LINK_END();

    STAMP(7);
    // ================= weight gradients (owner CTAs) -> Adam =================
    // D[input feature i (accumulator lane), output row j] = sum_b act[b, i] * dZ[b, j]: the activation
    // boxes read MN-major give M = i, 16 columns of dZ give N = 16, K = the batch rows.
    const bool last = t + 1 == a.n_steps;
    if (issuer) {
      if (lane == 0) {
        const uint32_t idesc = make_idesc(16, 3);
        for (int tt = 0; tt < 7; ++tt)
          for (int s = 0; s < 2; ++s)
            umma_bf16(tmem + cWg1 + 16 * tt, mn_desc(sb + oXB + 2 * tt * kBoxX, kBoxX) + (uint64_t)s * (2048 >> 4),
                      mn_desc(sb + oDZ1S, kBoxX) + (uint64_t)s * (2048 >> 4), idesc, s != 0 ? 1u : 0u);
        // rows 8c..8c+7 of W2: the 16 columns of dZ2 that contain them (slices start on 32-byte
        // boundaries inside the swizzled rows; CTAs 2k and 2k+1 read the same one and keep one half each)
        const uint32_t b2sl = sb + oDZ2 + (c >> 3) * kBoxX + ((c >> 1) & 3u) * 32u;
        for (int tt = 0; tt < 2; ++tt)
          for (int s = 0; s < 2; ++s)
            umma_bf16(tmem + cWg2 + 16 * tt, mn_desc(sb + oH1 + 2 * tt * kBoxX, kBoxX) + (uint64_t)s * (2048 >> 4),
                      mn_desc(b2sl, kBoxX) + (uint64_t)s * (2048 >> 4), idesc, s != 0 ? 1u : 0u);
        if (c == (uint32_t)kW3Owner)
          for (int s = 0; s < 2; ++s)
            umma_bf16(tmem + cWg3, mn_desc(sb + oH2, kBoxX) + (uint64_t)s * (2048 >> 4),
                      mn_desc(sb + oDZ3, kBoxX) + (uint64_t)s * (2048 >> 4), idesc, s != 0 ? 1u : 0u);
        umma_commit(smem_u32(&mb_mma));
      }
      __syncwarp();
    }
    // What Adam reads for W2 besides the gradient does not depend on this step: fetch it while the MMAs run
    float p2[kR2], m2[kR2], v2[kR2];
    const uint32_t w2base = kW2Off + (kR2 * c) * D1 + 128u * (g & 1) + 32u * q + lane;
    if (g < 2) {
#pragma unroll
      for (int j = 0; j < kR2; ++j) {
        p2[j] = __ldcg(a.flat + w2base + j * D1);
        m2[j] = __ldcg(a.m + w2base + j * D1);
        v2[j] = __ldcg(a.v + w2base + j * D1);
      }
    }
    LINK_WAIT();
    STAMP(8);
    if (one && !last) {
      // the batch buffer is free: the next batch travels while Adam runs
      mbar_expect_tx(smem_u32(&mb_x), kK0Boxes * kBoxX);
      for (int j = 0; j < kK0Boxes; ++j) tma_load_2d(&tmX, smem_u32(&mb_x), sb + oXB + j * kBoxX, 64 * j, row0 + kRows);
    }
    // --- W2 rows 8c..8c+7 first (the other CTAs wait for these): 2 tiles of 128 input features
    if (g < 2) {
      uint32_t r[16];
      tmem_ld16(tmem + lane_base + cWg2 + 16 * g, r);
#pragma unroll
      for (int j = 0; j < kR2; ++j) {
        const float gj = __uint_as_float((c & 1u) ? r[8 + j] : r[j]);
        adam_one(ak, gj, p2[j], m2[j], v2[j]);
        a.flat[w2base + j * D1] = p2[j];
        a.m[w2base + j * D1] = m2[j];
        a.v[w2base + j * D1] = v2[j];
        a.w16[w2base + j * D1] = __float2bfloat16_rn(p2[j]);
        if (a.grads) a.grads[w2base + j * D1] = gj;
      }
    }
    // --- W3 (CTA 8): one tile, 10 rows
    if (c == (uint32_t)kW3Owner && g == 2) {
      uint32_t r[16];
      tmem_ld16(tmem + lane_base + cWg3, r);
      const uint32_t i = 32u * q + lane;
      float p[16];
      adam_column<D3>(a, ak, r, kW3Off + i, D2, p);
    }
    // --- biases: the lanes that hold the batch sums
    if (warp == 0 && lane < kR1) {
      s_b1[lane] = adam_bias(a, ak, kB1Off + kR1 * c + lane, g_b1);
    }
    if (warp == 2 && lane < kR2) {
      adam_bias(a, ak, kB2Off + kR2 * c + lane, (s_db2[0][lane] + s_db2[1][lane]) + (s_db2[2][lane] + s_db2[3][lane]));
    }
    if (warp == 0 && c == (uint32_t)kW3Owner && lane < D3) adam_bias(a, ak, kB3Off + lane, g_b3);
    // --- W1 rows of this CTA: 7 tiles of 128 input features, two tiles per warp group; the moments
    // live in TMEM beside the gradient
#pragma unroll
    for (int sl = 0; sl < 2; ++sl) {
      const int tt = g + 4 * sl;
      if (tt < 7) {  // warp-uniform
        uint32_t r[16], rm[16], rv[16], rp[16];
        tmem_ld16(tmem + lane_base + cWg1 + 16 * tt, r);
        tmem_ld16(tmem + lane_base + cM + 16 * tt, rm);
        tmem_ld16(tmem + lane_base + cV + 16 * tt, rv);
        tmem_ld16(tmem + lane_base + cP + 16 * tt, rp);
        const uint32_t i = 128u * tt + 32u * q + lane;
        const bool valid = i < (uint32_t)D0;
        const uint32_t sdst = sb + oW1 + (i >> 6) * kBoxW1 + (((i & 63u) >> 3) << 4) + ((i & 7u) << 1);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          float pj = __uint_as_float(rp[j]), mj = __uint_as_float(rm[j]), vj = __uint_as_float(rv[j]);
          if (!valid) r[j] = 0u;
          adam_one(ak, __uint_as_float(r[j]), pj, mj, vj);
          rp[j] = __float_as_uint(pj);
          rm[j] = __float_as_uint(mj);
          rv[j] = __float_as_uint(vj);
          // this CTA's resident bf16 rows: row j, column i of the K-major swizzled boxes
          if (valid) st_shared_b16(sdst + j * 128u ^ ((uint32_t)(j & 7) << 4), __float2bfloat16_rn(pj));
        }
        tmem_st16(tmem + lane_base + cM + 16 * tt, rm);
        tmem_st16(tmem + lane_base + cV + 16 * tt, rv);
        tmem_st16(tmem + lane_base + cP + 16 * tt, rp);
        if (valid && (last || a.grads)) {
          // the epoch ends: master, moments and the bf16 copy go back to global memory
          const uint32_t idx0 = kW1Off + (kR1 * c) * D0 + i;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            if (last) {
              a.flat[idx0 + j * D0] = __uint_as_float(rp[j]);
              a.w16[idx0 + j * D0] = __float2bfloat16_rn(__uint_as_float(rp[j]));
              a.m[idx0 + j * D0] = __uint_as_float(rm[j]);
              a.v[idx0 + j * D0] = __uint_as_float(rv[j]);
            }
            if (a.grads) a.grads[idx0 + j * D0] = __uint_as_float(r[j]);
          }
        }
      }
    }
    tmem_st_wait();
    STAMP(9);
    fence_async_smem();      // the rewritten W1 rows vs. the next step's layer-1 MMAs
    fence_async_global();    // W2 / W3 in global memory vs. the other CTAs' TMA reloads (the barrier releases them)
    par ^= 1u;
    tc_fence_before();
    __syncthreads();         // this CTA's rows of W1 are rewritten, its accumulators are drained
    tc_fence_after();
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");  // (b), first half
    if (one && !last) issue_layer1(t + 1);  // issued and run while the other CTAs arrive
    __syncwarp();
    STAMP(10);
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");  // (b) every owner has published its rows
    tc_fence_after();
    STAMP(11);
    if (one && t + 1 < a.n_steps) {
      fence_async_global();
      mbar_expect_tx(smem_u32(&mb_w), 4 * kBoxW2 + 2 * kBoxW3);
      for (int j = 0; j < 4; ++j) tma_load_2d(&tmW2, smem_u32(&mb_w), sb + oW2 + j * kBoxW2, 64 * j, 0);
      for (int j = 0; j < 2; ++j) tma_load_2d(&tmW3, smem_u32(&mb_w), sb + oW3 + j * kBoxW3, 64 * j, 0);
    }
  }
#undef LINK_WAIT
#undef LINK_END
#undef STAMP
#undef STAMP1

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// bf16 [rows, ld] row-major, logical width `cols`; box = [64 columns x box_rows], 128B swizzle, zero fill
int encode(CUtensorMap* map, const void* base, int rows, int cols, int ld, int box_rows) {
  static EncodeTiledFn enc = nullptr;
  if (!enc) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      enc = reinterpret_cast<EncodeTiledFn>(p);
  }
  P2PFL_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%d cols=%d ld=%d box_rows=%d)", (int)r, rows, cols, ld,
              box_rows);
    return -2;
  }
  return 0;
}

int prepare(int device) {
  static bool done[64] = {false};
  if (device < 0 || device >= 64 || done[device]) return 0;
  P2PFL_CUDA_OK(cudaFuncSetAttribute(mlp_epoch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes + 1024));
  P2PFL_CUDA_OK(cudaFuncSetAttribute(mlp_epoch_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  done[device] = true;
  return 0;
}

void launch_config(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attr, cudaStream_t stream) {
  cfg = {};
  cfg.gridDim = dim3(kC);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = kSmemBytes + 1024;
  cfg.stream = stream;
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kC;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
}

}  // namespace

#ifdef P2PFL_BRINGUP
static long long* g_mlp_stamps = nullptr;
#endif
// Bring-up builds only (-DP2PFL_BRINGUP): `device_buffer` (int64 [16][32], NULL = off) receives the
// SM-clock stamps every CTA's warp 0 takes at the phase boundaries of one steady-state step.
extern "C" int p2pfl_mlp_epoch_set_stamps(void* device_buffer) {
#ifdef P2PFL_BRINGUP
  g_mlp_stamps = reinterpret_cast<long long*>(device_buffer);
  return 0;
#else
  if (device_buffer == nullptr) return 0;
  set_error("clock stamps are compiled out of this build (rebuild with -DP2PFL_BRINGUP)");
  return -1;
#endif
}

// 1 when this device can co-schedule the 16-CTA cluster the epoch kernel needs, else 0.
extern "C" int p2pfl_mlp_epoch_supported(int device) {
  if (cudaSetDevice(device) != cudaSuccess) return 0;
  if (prepare(device) != 0) {
    (void)cudaGetLastError();
    return 0;
  }
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[1];
  launch_config(cfg, attr, nullptr);
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, mlp_epoch_kernel, &cfg) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0;
  }
  return n >= 1 ? 1 : 0;
}

extern "C" int p2pfl_mlp_epoch(int device, void* stream, const p2pfl_mlp_epoch_args* e) {
  P2PFL_REQUIRE(e && e->xs && e->ys && e->flat && e->m && e->v && e->w16, "null argument");
  P2PFL_REQUIRE(e->n_rows > 0 && e->n_steps > 0 && (int64_t)e->n_steps * kRows >= e->n_rows &&
                    (int64_t)(e->n_steps - 1) * kRows < e->n_rows,
                "n_steps must be ceil(n_rows / 32)");
  P2PFL_REQUIRE((reinterpret_cast<uintptr_t>(e->xs) & 15) == 0 && (reinterpret_cast<uintptr_t>(e->w16) & 15) == 0,
                "xs and w16 must be 16-byte aligned");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  if (int rc = prepare(device)) return rc;
  CUtensorMap tmX, tmW1, tmW2, tmW3;
  const __nv_bfloat16* w16 = reinterpret_cast<const __nv_bfloat16*>(e->w16);
  // rows past the end of a tensor read as zeros: the last (partial) batch, W3's rows 10..15, columns 784..831
  if (int rc = encode(&tmX, e->xs, e->n_rows, D0, D0, kRows)) return rc;
  if (int rc = encode(&tmW1, w16 + kW1Off, D1, D0, D0, kR1)) return rc;
  if (int rc = encode(&tmW2, w16 + kW2Off, D2, D1, D1, D2)) return rc;
  if (int rc = encode(&tmW3, w16 + kW3Off, D3, D2, D2, 16)) return rc;
  Args a;
  a.ys = e->ys;
  a.flat = e->flat;
  a.m = e->m;
  a.v = e->v;
  a.w16 = reinterpret_cast<__nv_bfloat16*>(e->w16);
  a.grads = e->grads_out;
  a.loss = e->loss;
  a.step_ptr = e->step_ptr;
  a.n_rows = e->n_rows;
  a.n_steps = e->n_steps;
  a.lr = e->lr;
  a.b1 = e->beta1;
  a.b2 = e->beta2;
  a.eps = e->eps;
#ifdef P2PFL_BRINGUP
  a.stamps = g_mlp_stamps;
#endif
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[1];
  launch_config(cfg, attr, as_stream(stream));
  P2PFL_CUDA_OK(cudaLaunchKernelEx(&cfg, mlp_epoch_kernel, tmX, tmW1, tmW2, tmW3, a));
  return 0;
}
