// Fused softmax-cross-entropy forward+backward and the evaluation reduce.
// One warp per sample row; classes are strided over the lanes, so C = 10
// (MNIST/CIFAR) and C in the thousands both work.
#include "common.cuh"

using namespace p2pfl;

namespace {

// reference: mlp.py:68 (log_softmax) + mlp.py:78 (F.cross_entropy) + autograd;
// cnn.py:73,96 (CrossEntropyLoss).  dlogits = (softmax - onehot) / B.
__global__ void __launch_bounds__(256)
softmax_xent_kernel(const float* __restrict__ logits, int ld_logits, const int32_t* __restrict__ labels, int B, int C,
                    __nv_bfloat16* __restrict__ dl, int ld_d, __nv_bfloat16* __restrict__ dlt, int ld_dt,
                    float* __restrict__ dbias, float* __restrict__ loss_sum, int single_block) {
  // single_block: one CTA owns every row, so the bias gradient is reduced in shared
  // memory and STORED (the training step's case, B <= 256); otherwise it is
  // accumulated with atomics into a buffer the host zeroed
  __shared__ float s_db[1024];
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int warps = blockDim.x >> 5;
  const float inv_b = 1.f / (float)B;
  float block_loss = 0.f;
  if (single_block) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) s_db[c] = 0.f;
    __syncthreads();
  }
  for (int row = blockIdx.x * warps + warp; row < B; row += gridDim.x * warps) {
    const float* x = logits + (uint64_t)row * ld_logits;
    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) mx = fmaxf(mx, x[c]);
    mx = warp_max(mx);
    float se = 0.f;
    for (int c = lane; c < C; c += 32) se += __expf(x[c] - mx);
    se = warp_sum(se);
    const float lse = mx + __logf(se);
    const int y = labels[row];
    for (int c = lane; c < C; c += 32) {
      float p = __expf(x[c] - lse);
      float g = (p - (c == y ? 1.f : 0.f)) * inv_b;
      __nv_bfloat16 gb = __float2bfloat16_rn(g);
      if (dl) dl[(uint64_t)row * ld_d + c] = gb;
      if (dlt) dlt[(uint64_t)c * ld_dt + row] = gb;
      // the bias gradient is the column sum of exactly what the GEMMs will see
      if (dbias) atomicAdd(single_block ? s_db + c : dbias + c, __bfloat162float(gb));
    }
    if (lane == 0 && y >= 0 && y < C) block_loss += lse - x[y];
  }
  if (single_block && dbias) {
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) dbias[c] = s_db[c];
  }
  if (loss_sum) {
    __shared__ float sh[8];
    if (lane == 0) sh[warp] = block_loss;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int w = 0; w < warps; ++w) t += sh[w];
      if (t != 0.f) atomicAdd(loss_sum, t);
    }
  }
}

// reference: mlp.py:96-99 / cnn.py:114-117 (test_step: loss + torchmetrics Accuracy)
__global__ void __launch_bounds__(256)
eval_reduce_kernel(const float* __restrict__ logits, int ld_logits, const int32_t* __restrict__ labels, int B, int C,
                   float* __restrict__ loss_sum, int32_t* __restrict__ correct) {
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int warps = blockDim.x >> 5;
  float w_loss = 0.f;
  int w_correct = 0;
  for (int row = blockIdx.x * warps + warp; row < B; row += gridDim.x * warps) {
    const float* x = logits + (uint64_t)row * ld_logits;
    float mx = -INFINITY;
    int arg = 0;
    for (int c = lane; c < C; c += 32) {
      float v = x[c];
      if (v > mx) {
        mx = v;
        arg = c;
      }
    }
    // warp argmax (first index wins ties, like torch.argmax)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float om = __shfl_xor_sync(0xffffffffu, mx, o);
      int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > mx || (om == mx && oa < arg)) {
        mx = om;
        arg = oa;
      }
    }
    float se = 0.f;
    for (int c = lane; c < C; c += 32) se += __expf(x[c] - mx);
    se = warp_sum(se);
    const int y = labels[row];
    if (lane == 0) {
      w_loss += mx + __logf(se) - x[y];
      w_correct += (arg == y) ? 1 : 0;
    }
  }
  __shared__ float shl[8];
  __shared__ int shc[8];
  if (lane == 0) {
    shl[warp] = w_loss;
    shc[warp] = w_correct;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float tl = 0.f;
    int tc = 0;
    for (int w = 0; w < warps; ++w) {
      tl += shl[w];
      tc += shc[w];
    }
    if (loss_sum) atomicAdd(loss_sum, tl);
    if (correct) atomicAdd(correct, tc);
  }
}

}  // namespace

extern "C" {

int p2pfl_softmax_xent(int device, void* stream, const float* logits, int ld_logits, const int32_t* labels, int B, int C,
                       void* dlogits_bf16, int ld_d, void* dlogits_t_bf16, int ld_dt, float* dbias, float* loss_sum) {
  P2PFL_REQUIRE(logits && labels && B > 0 && C > 0, "bad arguments to p2pfl_softmax_xent");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  const int single = (B <= 256 && C <= 1024) ? 1 : 0;
  int blocks = single ? 1 : (B + 7) / 8;
  int cap = sm_count(device) * 4;
  if (!single && dbias) P2PFL_CUDA_OK(cudaMemsetAsync(dbias, 0, sizeof(float) * C, as_stream(stream)));
  P2PFL_CUDA_OK(launch_pdl(softmax_xent_kernel, dim3(blocks < cap ? blocks : cap), dim3(256), 0, as_stream(stream), logits,
                           ld_logits, labels, B, C, (__nv_bfloat16*)dlogits_bf16, ld_d, (__nv_bfloat16*)dlogits_t_bf16,
                           ld_dt, dbias, loss_sum, single));
  return 0;
}

int p2pfl_eval_reduce(int device, void* stream, const float* logits, int ld_logits, const int32_t* labels, int B, int C,
                      float* loss_sum, int32_t* correct) {
  P2PFL_REQUIRE(logits && labels && B > 0 && C > 0, "bad arguments to p2pfl_eval_reduce");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  int blocks = (B + 7) / 8;
  int cap = sm_count(device) * 4;
  eval_reduce_kernel<<<blocks < cap ? blocks : cap, 256, 0, as_stream(stream)>>>(logits, ld_logits, labels, B, C,
                                                                                 loss_sum, correct);
  P2PFL_LAUNCH_OK();
  return 0;
}

}  // extern "C"
