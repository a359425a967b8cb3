// Memory-bound fused kernels over the flat parameter arena: Adam / SGD update
// (+ bf16 working copy), FedAvg weighted sum, casts, transposes, batch gather.
// All are single-pass, 128-bit vectorised, grid sized to the SM count.
#include "common.cuh"

using namespace p2pfl;

namespace {

constexpr int kThreads = 256;

inline int grid_for(uint64_t work_items, int device, int per_sm = 8) {
  uint64_t blocks = (work_items + kThreads - 1) / kThreads;
  uint64_t cap = (uint64_t)sm_count(device) * per_sm;
  if (blocks < 1) blocks = 1;
  return (int)(blocks < cap ? blocks : cap);
}

struct TransposeBatch {
  p2pfl_transpose_desc d[16];
  int32_t tile_start[17];  // prefix sum of 32x32 tiles per matrix
  int32_t count;
};

// write element `idx` of the flat fp32 arena into every bf16 transposed copy that covers it
__device__ __forceinline__ void scatter_transposed(const TransposeBatch& tb, __nv_bfloat16* __restrict__ wt, uint64_t idx,
                                                   float v) {
  for (int k = 0; k < tb.count; ++k) {
    const p2pfl_transpose_desc& d = tb.d[k];
    const uint64_t rel = idx - d.src_off;
    if (idx >= d.src_off && rel < (uint64_t)d.rows * d.cols) {
      const uint32_t r = (uint32_t)(rel / (uint32_t)d.cols), c = (uint32_t)(rel - (uint64_t)r * d.cols);
      wt[d.dst_off + (uint64_t)c * d.ld_t + r] = __float2bfloat16_rn(v);
    }
  }
}

// --------------------------------------------------------------------- Adam
// reference: torch.optim.Adam defaults as used by mlp.py:71-73 / cnn.py:89-91
__global__ void __launch_bounds__(kThreads)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
            __nv_bfloat16* __restrict__ p16, uint64_t n, float lr, float b1, float b2, float eps, float wd,
            const int32_t* __restrict__ step_ptr, int32_t step_offset, __nv_bfloat16* __restrict__ wt,
            const __grid_constant__ TransposeBatch tb) {
  pdl_launch_dependents();
  pdl_wait();
  const int t = (step_ptr ? *step_ptr : 0) + step_offset;
  const float bc1 = 1.f - powf(b1, (float)t);
  const float bc2 = 1.f - powf(b2, (float)t);
  const float step_size = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  const uint64_t n4 = n >> 2;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    float4 gv = reinterpret_cast<const float4*>(g)[i];
    float4 mv = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float pe[4] = {pv.x, pv.y, pv.z, pv.w}, ge[4] = {gv.x, gv.y, gv.z, gv.w};
    float me[4] = {mv.x, mv.y, mv.z, mv.w}, ve[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float gk = ge[k] + wd * pe[k];
      me[k] = b1 * me[k] + (1.f - b1) * gk;
      ve[k] = b2 * ve[k] + (1.f - b2) * gk * gk;
      float denom = sqrtf(ve[k]) * inv_sqrt_bc2 + eps;
      pe[k] -= step_size * (me[k] / denom);
    }
    reinterpret_cast<float4*>(p)[i] = make_float4(pe[0], pe[1], pe[2], pe[3]);
    reinterpret_cast<float4*>(m)[i] = make_float4(me[0], me[1], me[2], me[3]);
    reinterpret_cast<float4*>(v)[i] = make_float4(ve[0], ve[1], ve[2], ve[3]);
    if (p16) reinterpret_cast<uint2*>(p16)[i] = make_uint2(pack_bf16x2(pe[0], pe[1]), pack_bf16x2(pe[2], pe[3]));
    if (wt) {
#pragma unroll
      for (int k = 0; k < 4; ++k) scatter_transposed(tb, wt, i * 4 + k, pe[k]);
    }
  }
  // tail (n not a multiple of 4)
  for (uint64_t i = (n4 << 2) + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float gk = g[i] + wd * p[i];
    float mk = b1 * m[i] + (1.f - b1) * gk;
    float vk = b2 * v[i] + (1.f - b2) * gk * gk;
    float pk = p[i] - step_size * (mk / (sqrtf(vk) * inv_sqrt_bc2 + eps));
    p[i] = pk;
    m[i] = mk;
    v[i] = vk;
    if (p16) p16[i] = __float2bfloat16_rn(pk);
    if (wt) scatter_transposed(tb, wt, i, pk);
  }
}

__global__ void __launch_bounds__(kThreads)
sgd_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ mom,
           __nv_bfloat16* __restrict__ p16, uint64_t n, float lr, float momentum) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float d = g[i];
    if (mom) {
      d = momentum * mom[i] + d;
      mom[i] = d;
    }
    float pk = p[i] - lr * d;
    p[i] = pk;
    if (p16) p16[i] = __float2bfloat16_rn(pk);
  }
}

__global__ void add_i32_kernel(int32_t* ptr, int32_t delta) { *ptr += delta; }

__global__ void __launch_bounds__(kThreads)
cast_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, uint64_t n) {
  const uint64_t n4 = n >> 2;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 v = reinterpret_cast<const float4*>(src)[i];
    reinterpret_cast<uint2*>(dst)[i] = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  }
  for (uint64_t i = (n4 << 2) + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = __float2bfloat16_rn(src[i]);
}

// ------------------------------------------------------- batched transpose
__global__ void __launch_bounds__(256)
transpose_cast_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, TransposeBatch tb) {
  __shared__ float tile[32][33];
  for (int tidx = blockIdx.x; tidx < tb.tile_start[tb.count]; tidx += gridDim.x) {
    int mi = 0;
    while (tidx >= tb.tile_start[mi + 1]) ++mi;
    const p2pfl_transpose_desc d = tb.d[mi];
    const int local = tidx - tb.tile_start[mi];
    const int tiles_c = (d.cols + 31) / 32;
    const int tr = local / tiles_c, tc = local % tiles_c;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    const float* s = src + d.src_off;
    __nv_bfloat16* o = dst + d.dst_off;
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      int r = tr * 32 + ty + j, c = tc * 32 + tx;
      tile[ty + j][tx] = (r < d.rows && c < d.cols) ? s[(uint64_t)r * d.cols + c] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      int c = tc * 32 + ty + j, r = tr * 32 + tx;  // output row = c, col = r
      if (c < d.cols && r < d.rows) o[(uint64_t)c * d.ld_t + r] = __float2bfloat16_rn(tile[tx][ty + j]);
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------ FedAvg
struct FedAvgArgs {
  const void* model[P2PFL_MAX_MODELS];
  float weight[P2PFL_MAX_MODELS];
  int32_t is_bf16[P2PFL_MAX_MODELS];
  int32_t n_models;
};

// reference: fedavg.py:49-58 (accum += m_i * w_i ; accum /= sum w) - weights arrive normalised
__global__ void __launch_bounds__(kThreads)
fedavg_kernel(FedAvgArgs a, float* __restrict__ out_f32, __nv_bfloat16* __restrict__ out_bf16, uint64_t n) {
  const uint64_t n4 = n >> 2;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
    for (int k = 0; k < a.n_models; ++k) {
      const float w = a.weight[k];
      if (a.is_bf16[k]) {
        uint2 u = reinterpret_cast<const uint2*>(a.model[k])[i];
        float2 lo = unpack_bf16x2(u.x), hi = unpack_bf16x2(u.y);
        acc[0] = fmaf(w, lo.x, acc[0]);
        acc[1] = fmaf(w, lo.y, acc[1]);
        acc[2] = fmaf(w, hi.x, acc[2]);
        acc[3] = fmaf(w, hi.y, acc[3]);
      } else {
        float4 v = reinterpret_cast<const float4*>(a.model[k])[i];
        acc[0] = fmaf(w, v.x, acc[0]);
        acc[1] = fmaf(w, v.y, acc[1]);
        acc[2] = fmaf(w, v.z, acc[2]);
        acc[3] = fmaf(w, v.w, acc[3]);
      }
    }
    if (out_f32) reinterpret_cast<float4*>(out_f32)[i] = make_float4(acc[0], acc[1], acc[2], acc[3]);
    if (out_bf16)
      reinterpret_cast<uint2*>(out_bf16)[i] = make_uint2(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]));
  }
  for (uint64_t i = (n4 << 2) + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float acc = 0.f;
    for (int k = 0; k < a.n_models; ++k) {
      float x = a.is_bf16[k] ? bf16_to_f32(reinterpret_cast<const __nv_bfloat16*>(a.model[k])[i])
                             : reinterpret_cast<const float*>(a.model[k])[i];
      acc = fmaf(a.weight[k], x, acc);
    }
    if (out_f32) out_f32[i] = acc;
    if (out_bf16) out_bf16[i] = __float2bfloat16_rn(acc);
  }
}

// -------------------------------------------------------------- row gather
// One warp per output row: dst[i] = bf16(src[perm[i]]); optional per-batch
// transposed copy dst_t[i / B][f][i % B] feeding the wgrad GEMM (K = batch).
__global__ void __launch_bounds__(256)
gather_rows_kernel(const float* __restrict__ src, const int32_t* __restrict__ perm, int n_rows, int feat, int batch,
                   __nv_bfloat16* __restrict__ dst, __nv_bfloat16* __restrict__ dst_t,
                   const int32_t* __restrict__ labels_src, int32_t* __restrict__ labels_dst) {
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  for (int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < n_rows; row += gridDim.x * warps_per_block) {
    const int srow = perm ? perm[row] : row;
    const float* s = src + (uint64_t)srow * feat;
    __nv_bfloat16* d = dst + (uint64_t)row * feat;
    __nv_bfloat16* dt = dst_t ? dst_t + (uint64_t)(row / batch) * feat * batch + (row % batch) : nullptr;
    for (int f = lane; f < feat; f += 32) {
      __nv_bfloat16 v = __float2bfloat16_rn(s[f]);
      d[f] = v;
      if (dt) dt[(uint64_t)f * batch] = v;
    }
    if (lane == 0 && labels_src && labels_dst) labels_dst[row] = labels_src[srow];
  }
}

}  // namespace

extern "C" {

int p2pfl_adam_step(int device, void* stream, float* p, const float* g, float* m, float* v, void* p_bf16, uint64_t n,
                    float lr, float beta1, float beta2, float eps, float weight_decay, const int32_t* step_ptr,
                    int32_t step_offset) {
  P2PFL_REQUIRE(p && g && m && v, "null pointer passed to p2pfl_adam_step");
  P2PFL_REQUIRE((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0, "adam buffers must be 16-byte aligned");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  TransposeBatch tb;
  tb.count = 0;
  P2PFL_CUDA_OK(launch_pdl(adam_kernel, dim3(grid_for(n / 4 + 1, device)), dim3(kThreads), 0, as_stream(stream), p, g, m, v,
                           (__nv_bfloat16*)p_bf16, n, lr, beta1, beta2, eps, weight_decay, step_ptr, step_offset,
                           (__nv_bfloat16*)nullptr, tb));
  return 0;
}

int p2pfl_adam_step_t(int device, void* stream, float* p, const float* g, float* m, float* v, void* p_bf16, uint64_t n,
                      float lr, float beta1, float beta2, float eps, float weight_decay, const int32_t* step_ptr,
                      int32_t step_offset, void* wt_bf16, const p2pfl_transpose_desc* descs, int count) {
  P2PFL_REQUIRE(p && g && m && v, "null pointer passed to p2pfl_adam_step_t");
  P2PFL_REQUIRE((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0, "adam buffers must be 16-byte aligned");
  P2PFL_REQUIRE(count >= 0 && count <= 16 && (count == 0 || (wt_bf16 && descs)), "bad transpose list");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  TransposeBatch tb;
  tb.count = count;
  for (int i = 0; i < count; ++i) tb.d[i] = descs[i];
  P2PFL_CUDA_OK(launch_pdl(adam_kernel, dim3(grid_for(n / 4 + 1, device)), dim3(kThreads), 0, as_stream(stream), p, g, m, v,
                           (__nv_bfloat16*)p_bf16, n, lr, beta1, beta2, eps, weight_decay, step_ptr, step_offset,
                           count ? (__nv_bfloat16*)wt_bf16 : (__nv_bfloat16*)nullptr, tb));
  return 0;
}

int p2pfl_sgd_step(int device, void* stream, float* p, const float* g, float* mom, void* p_bf16, uint64_t n, float lr,
                   float momentum) {
  P2PFL_REQUIRE(p && g, "null pointer passed to p2pfl_sgd_step");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  sgd_kernel<<<grid_for(n, device), kThreads, 0, as_stream(stream)>>>(p, g, mom, (__nv_bfloat16*)p_bf16, n, lr, momentum);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_add_i32(int device, void* stream, int32_t* ptr, int32_t delta) {
  P2PFL_CUDA_OK(cudaSetDevice(device));
  add_i32_kernel<<<1, 1, 0, as_stream(stream)>>>(ptr, delta);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_cast_f32_bf16(int device, void* stream, const float* src, void* dst, uint64_t n) {
  P2PFL_REQUIRE((((uintptr_t)src & 15) | ((uintptr_t)dst & 7)) == 0, "cast buffers must be 16/8-byte aligned");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  cast_kernel<<<grid_for(n / 4 + 1, device), kThreads, 0, as_stream(stream)>>>(src, (__nv_bfloat16*)dst, n);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_transpose_cast(int device, void* stream, const float* src, void* dst, const p2pfl_transpose_desc* descs,
                         int count) {
  P2PFL_REQUIRE(count >= 1 && count <= 16, "p2pfl_transpose_cast handles 1..16 matrices per call");
  TransposeBatch tb;
  tb.count = count;
  tb.tile_start[0] = 0;
  for (int i = 0; i < count; ++i) {
    tb.d[i] = descs[i];
    tb.tile_start[i + 1] = tb.tile_start[i] + ((descs[i].rows + 31) / 32) * ((descs[i].cols + 31) / 32);
  }
  P2PFL_CUDA_OK(cudaSetDevice(device));
  int total = tb.tile_start[count];
  int cap = sm_count(device) * 8;
  transpose_cast_kernel<<<total < cap ? total : cap, 256, 0, as_stream(stream)>>>(src, (__nv_bfloat16*)dst, tb);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_fedavg(int device, void* stream, int n_models, const void* const* models, const float* weights,
                 const int32_t* is_bf16, float* out_f32, void* out_bf16, uint64_t n) {
  P2PFL_REQUIRE(n_models >= 1 && n_models <= P2PFL_MAX_MODELS, "p2pfl_fedavg takes 1..16 models per call");
  P2PFL_REQUIRE(out_f32 || out_bf16, "p2pfl_fedavg needs an output");
  FedAvgArgs a;
  a.n_models = n_models;
  for (int i = 0; i < n_models; ++i) {
    P2PFL_REQUIRE(((uintptr_t)models[i] & 15) == 0, "fedavg inputs must be 16-byte aligned");
    a.model[i] = models[i];
    a.weight[i] = weights[i];
    a.is_bf16[i] = is_bf16 ? is_bf16[i] : 0;
  }
  P2PFL_CUDA_OK(cudaSetDevice(device));
  fedavg_kernel<<<grid_for(n / 4 + 1, device), kThreads, 0, as_stream(stream)>>>(a, out_f32, (__nv_bfloat16*)out_bf16, n);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_gather_rows(int device, void* stream, const float* src, const int32_t* perm, int n_rows, int feat, int batch,
                      void* dst_bf16, void* dst_t_bf16, const int32_t* labels_src, int32_t* labels_dst) {
  P2PFL_REQUIRE(src && dst_bf16 && batch > 0, "bad arguments to p2pfl_gather_rows");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  int blocks = (n_rows + 7) / 8;
  int cap = sm_count(device) * 8;
  gather_rows_kernel<<<blocks < cap ? blocks : cap, 256, 0, as_stream(stream)>>>(
      src, perm, n_rows, feat, batch, (__nv_bfloat16*)dst_bf16, (__nv_bfloat16*)dst_t_bf16, labels_src, labels_dst);
  P2PFL_LAUNCH_OK();
  return 0;
}

}  // extern "C"
