// Convolution-path kernels for the CNN / ResNet families.
//
// reference sites replaced: cnn.py:78-81 (Conv2d 5x5 "same" + ReLU + MaxPool2d via
// cuDNN / ATen) and their autograd; ResNet-style blocks (BatchNorm, residual add,
// global average pool) for the BASELINE.json ResNet-18 config.
//
// Convolutions run as implicit GEMMs on the tcgen05 kernel (gemm_tcgen05.cu): the
// kernels here build the GEMM operands in the orientations that kernel wants
// (activations are NHWC bf16; "col" = im2col matrix [pixels, Cin*kh*kw] with the
// K index ordered (c, i, j) exactly like a flattened PyTorch conv weight, plus its
// transpose for the weight-gradient GEMM whose K dimension is the pixel count), and
// fold everything that is not a GEMM - ReLU masks, pooling, bias gradients, layout
// changes - into single passes.
#include "common.cuh"

using namespace p2pfl;

namespace {

constexpr int kThreads = 256;

inline int grid_for(uint64_t items, int device) {
  uint64_t blocks = (items + kThreads - 1) / kThreads;
  uint64_t cap = (uint64_t)sm_count(device) * 16;
  if (blocks < 1) blocks = 1;
  return (int)(blocks < cap ? blocks : cap);
}

// ------------------------------------------------------------------- im2col
// x: NHWC bf16 [B,H,W,C].  col[(b,ho,wo), c*kh*kw + i*kw + j] = x[b, ho*s+i-p, wo*s+j-p, c]
// col_t (nullable): the transpose [K, ld_t] (ld_t >= pixels).
__global__ void __launch_bounds__(kThreads)
im2col_kernel(const __nv_bfloat16* __restrict__ x, int B, int H, int W, int C, int kh, int kw, int stride, int pad,
              int Ho, int Wo, __nv_bfloat16* __restrict__ col, int ld_col, __nv_bfloat16* __restrict__ col_t,
              int64_t ld_t) {
  const int K = C * kh * kw;
  const int64_t total = (int64_t)B * Ho * Wo * K;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int k = (int)(idx % K);
    const int64_t pix = idx / K;
    const int wo = (int)(pix % Wo);
    const int ho = (int)((pix / Wo) % Ho);
    const int b = (int)(pix / ((int64_t)Wo * Ho));
    const int c = k / (kh * kw);
    const int ij = k - c * kh * kw;
    const int i = ij / kw, j = ij - i * kw;
    const int h = ho * stride + i - pad, w = wo * stride + j - pad;
    __nv_bfloat16 v = __float2bfloat16_rn(0.f);
    if (h >= 0 && h < H && w >= 0 && w < W) v = x[(((int64_t)b * H + h) * W + w) * C + c];
    col[pix * ld_col + k] = v;
    if (col_t) col_t[(int64_t)k * ld_t + pix] = v;
  }
}

// ------------------------------------------------------------------- col2im
// dx[b,h,w,c] = sum over (i,j) of dcol[(b,ho,wo), c*kh*kw + i*kw + j] with ho*s+i-p == h, wo*s+j-p == w.
// Gather form: one thread per input element, no atomics.
__global__ void __launch_bounds__(kThreads)
col2im_kernel(const __nv_bfloat16* __restrict__ dcol, int ld_col, int B, int H, int W, int C, int kh, int kw, int stride,
              int pad, int Ho, int Wo, __nv_bfloat16* __restrict__ dx) {
  const int64_t total = (int64_t)B * H * W * C;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % C);
    const int w = (int)((idx / C) % W);
    const int h = (int)((idx / ((int64_t)C * W)) % H);
    const int b = (int)(idx / ((int64_t)C * W * H));
    float acc = 0.f;
    for (int i = 0; i < kh; ++i) {
      const int hn = h + pad - i;
      if (hn < 0 || hn % stride) continue;
      const int ho = hn / stride;
      if (ho >= Ho) continue;
      for (int j = 0; j < kw; ++j) {
        const int wn = w + pad - j;
        if (wn < 0 || wn % stride) continue;
        const int wo = wn / stride;
        if (wo >= Wo) continue;
        const int64_t pix = ((int64_t)b * Ho + ho) * Wo + wo;
        acc += __bfloat162float(dcol[pix * ld_col + c * kh * kw + i * kw + j]);
      }
    }
    dx[idx] = __float2bfloat16_rn(acc);
  }
}

// --------------------------------------------------------------- 2x2 max pool
// x NHWC [B,H,W,C] (H, W even) -> y [B,H/2,W/2,C], arg (0..3) per output.
// flat / flat_t (nullable): the same values in NCHW-flatten order [B, C*Ho*Wo] and its
// transpose [C*Ho*Wo, ld_flat_t] - what the following Linear layer's GEMMs consume.
__global__ void __launch_bounds__(kThreads)
maxpool2_fwd_kernel(const __nv_bfloat16* __restrict__ x, int B, int H, int W, int C, __nv_bfloat16* __restrict__ y,
                    uint8_t* __restrict__ arg, __nv_bfloat16* __restrict__ flat, __nv_bfloat16* __restrict__ flat_t,
                    int ld_flat_t) {
  const int Ho = H / 2, Wo = W / 2;
  const int64_t total = (int64_t)B * Ho * Wo * C;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % C);
    const int wo = (int)((idx / C) % Wo);
    const int ho = (int)((idx / ((int64_t)C * Wo)) % Ho);
    const int b = (int)(idx / ((int64_t)C * Wo * Ho));
    float best = -INFINITY;
    int a = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int h = ho * 2 + (q >> 1), w = wo * 2 + (q & 1);
      const float v = __bfloat162float(x[(((int64_t)b * H + h) * W + w) * C + c]);
      if (v > best) {
        best = v;
        a = q;
      }
    }
    const __nv_bfloat16 o = __float2bfloat16_rn(best);
    if (y) y[idx] = o;
    arg[idx] = (uint8_t)a;
    if (flat) {
      const int f = (c * Ho + ho) * Wo + wo;
      flat[(int64_t)b * (C * Ho * Wo) + f] = o;
      if (flat_t) flat_t[(int64_t)f * ld_flat_t + b] = o;
    }
  }
}

// Backward of (ReLU -> 2x2 max pool): routes dy to the argmax position, zeroes it where
// the pre-pool activation was not positive, and writes the result in both orientations
// ([pixels, C] and [C, ld_t]) plus the per-channel sum (bias gradient, atomics into dbias).
// dy comes either NHWC (dy_nhwc) or NCHW-flatten (dy_flat) - whichever is non-null.
__global__ void __launch_bounds__(kThreads)
maxpool2_relu_bwd_kernel(const __nv_bfloat16* __restrict__ dy_nhwc, const __nv_bfloat16* __restrict__ dy_flat,
                         const uint8_t* __restrict__ arg, const __nv_bfloat16* __restrict__ act, int B, int H, int W,
                         int C, __nv_bfloat16* __restrict__ dx, __nv_bfloat16* __restrict__ dx_t, int64_t ld_t,
                         float* __restrict__ dbias) {
  const int Ho = H / 2, Wo = W / 2;
  const int64_t total = (int64_t)B * H * W * C;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % C);
    const int w = (int)((idx / C) % W);
    const int h = (int)((idx / ((int64_t)C * W)) % H);
    const int b = (int)(idx / ((int64_t)C * W * H));
    const int ho = h >> 1, wo = w >> 1, q = ((h & 1) << 1) | (w & 1);
    const int64_t oidx = (((int64_t)b * Ho + ho) * Wo + wo) * C + c;
    float g = 0.f;
    if (arg[oidx] == q && __bfloat162float(act[idx]) > 0.f) {
      g = dy_nhwc ? __bfloat162float(dy_nhwc[oidx])
                  : __bfloat162float(dy_flat[(int64_t)b * (C * Ho * Wo) + (c * Ho + ho) * Wo + wo]);
    }
    const __nv_bfloat16 gb = __float2bfloat16_rn(g);
    dx[idx] = gb;
    const int64_t pix = idx / C;
    if (dx_t) dx_t[(int64_t)c * ld_t + pix] = gb;
    if (dbias && g != 0.f) atomicAdd(dbias + c, __bfloat162float(gb));
  }
}

// ------------------------------------------------------------- pad / cast copy
// rows x cols fp32 -> bf16 [rows, ld] (ld >= cols), padding zeroed: K-major GEMM operands
// need a 16-byte row pitch (first-layer conv weights have K = 25 or 27).
__global__ void __launch_bounds__(kThreads)
pad_cast_kernel(const float* __restrict__ src, int rows, int cols, __nv_bfloat16* __restrict__ dst, int ld) {
  const int64_t total = (int64_t)rows * ld;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % ld);
    const int64_t r = idx / ld;
    dst[idx] = __float2bfloat16_rn(c < cols ? src[r * cols + c] : 0.f);
  }
}

// NCHW fp32 -> NHWC bf16 (network input)
__global__ void __launch_bounds__(kThreads)
nchw_to_nhwc_kernel(const float* __restrict__ x, int B, int C, int H, int W, __nv_bfloat16* __restrict__ y) {
  const int64_t total = (int64_t)B * C * H * W;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % C);
    const int w = (int)((idx / C) % W);
    const int h = (int)((idx / ((int64_t)C * W)) % H);
    const int b = (int)(idx / ((int64_t)C * W * H));
    y[idx] = __float2bfloat16_rn(x[(((int64_t)b * C + c) * H + h) * W + w]);
  }
}

// ---------------------------------------------------------------- BatchNorm
// Training-mode BatchNorm2d over NHWC: one block per channel reduces mean / biased
// variance over the B*H*W pixels, updates the running statistics, normalises.
__device__ __forceinline__ float block_sum(float v, float* sh) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  float t = 0.f;
  if (warp == 0) {
    t = lane < (blockDim.x >> 5) ? sh[lane] : 0.f;
    t = warp_sum(t);
    if (lane == 0) sh[0] = t;
  }
  __syncthreads();
  t = sh[0];
  __syncthreads();
  return t;
}

__global__ void __launch_bounds__(kThreads)
bn_fwd_kernel(const __nv_bfloat16* __restrict__ x, int64_t pixels, int C, const float* __restrict__ gamma,
              const float* __restrict__ beta, float* __restrict__ running_mean, float* __restrict__ running_var,
              float momentum, float eps, int relu, const __nv_bfloat16* __restrict__ residual,
              __nv_bfloat16* __restrict__ y, float* __restrict__ save_mean, float* __restrict__ save_rstd) {
  __shared__ float sh[32];
  for (int c = blockIdx.x; c < C; c += gridDim.x) {
    float s = 0.f, ss = 0.f;
    for (int64_t p = threadIdx.x; p < pixels; p += blockDim.x) {
      const float v = __bfloat162float(x[p * C + c]);
      s += v;
      ss += v * v;
    }
    s = block_sum(s, sh);
    ss = block_sum(ss, sh);
    const float mean = s / (float)pixels;
    const float var = fmaxf(ss / (float)pixels - mean * mean, 0.f);
    const float rstd = rsqrtf(var + eps);
    if (threadIdx.x == 0) {
      save_mean[c] = mean;
      save_rstd[c] = rstd;
      if (running_mean) {
        const float unbiased = pixels > 1 ? var * (float)pixels / (float)(pixels - 1) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
      }
    }
    const float g = gamma[c], bt = beta[c];
    for (int64_t p = threadIdx.x; p < pixels; p += blockDim.x) {
      float v = (__bfloat162float(x[p * C + c]) - mean) * rstd * g + bt;
      if (residual) v += __bfloat162float(residual[p * C + c]);
      if (relu) v = fmaxf(v, 0.f);
      y[p * C + c] = __float2bfloat16_rn(v);
    }
  }
}

// Backward of y = [relu]( bn(x) [+ residual] ):  dy is masked by (y > 0) when relu,
// passed through to d_residual, and pushed through the normalisation.
__global__ void __launch_bounds__(kThreads)
bn_bwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ dy,
              int64_t pixels, int C, const float* __restrict__ gamma, const float* __restrict__ save_mean,
              const float* __restrict__ save_rstd, int relu, __nv_bfloat16* __restrict__ dx,
              __nv_bfloat16* __restrict__ d_residual, float* __restrict__ dgamma, float* __restrict__ dbeta) {
  __shared__ float sh[32];
  for (int c = blockIdx.x; c < C; c += gridDim.x) {
    const float mean = save_mean[c], rstd = save_rstd[c], g = gamma[c];
    float sdy = 0.f, sdyx = 0.f;
    for (int64_t p = threadIdx.x; p < pixels; p += blockDim.x) {
      float d = __bfloat162float(dy[p * C + c]);
      if (relu && !(__bfloat162float(y[p * C + c]) > 0.f)) d = 0.f;
      const float xh = (__bfloat162float(x[p * C + c]) - mean) * rstd;
      sdy += d;
      sdyx += d * xh;
    }
    sdy = block_sum(sdy, sh);
    sdyx = block_sum(sdyx, sh);
    if (threadIdx.x == 0) {
      dgamma[c] = sdyx;
      dbeta[c] = sdy;
    }
    const float inv_n = 1.f / (float)pixels;
    for (int64_t p = threadIdx.x; p < pixels; p += blockDim.x) {
      float d = __bfloat162float(dy[p * C + c]);
      if (relu && !(__bfloat162float(y[p * C + c]) > 0.f)) d = 0.f;
      if (d_residual) d_residual[p * C + c] = __float2bfloat16_rn(d);
      const float xh = (__bfloat162float(x[p * C + c]) - mean) * rstd;
      dx[p * C + c] = __float2bfloat16_rn(g * rstd * (d - sdy * inv_n - xh * sdyx * inv_n));
    }
  }
}

// ------------------------------------------------------ global average pool
// x NHWC [B, HW, C] -> y [B, C] (and optional transpose [C, ld_t]); backward broadcasts dy / HW.
__global__ void __launch_bounds__(kThreads)
avgpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, int B, int HW, int C, __nv_bfloat16* __restrict__ y,
                   __nv_bfloat16* __restrict__ y_t, int ld_t) {
  const int64_t total = (int64_t)B * C;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % C);
    const int b = (int)(idx / C);
    float s = 0.f;
    for (int p = 0; p < HW; ++p) s += __bfloat162float(x[((int64_t)b * HW + p) * C + c]);
    const __nv_bfloat16 o = __float2bfloat16_rn(s / (float)HW);
    y[idx] = o;
    if (y_t) y_t[(int64_t)c * ld_t + b] = o;
  }
}

__global__ void __launch_bounds__(kThreads)
avgpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, int B, int HW, int C, __nv_bfloat16* __restrict__ dx) {
  const int64_t total = (int64_t)B * HW * C;
  const float inv = 1.f / (float)HW;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % C);
    const int b = (int)(idx / ((int64_t)C * HW));
    dx[idx] = __float2bfloat16_rn(__bfloat162float(dy[(int64_t)b * C + c]) * inv);
  }
}

}  // namespace

extern "C" {

int p2pfl_im2col(int device, void* stream, const void* x_nhwc, int B, int H, int W, int C, int kh, int kw, int stride,
                 int pad, void* col, int ld_col, void* col_t, int64_t ld_t) {
  P2PFL_REQUIRE(x_nhwc && col && B > 0 && C > 0 && kh > 0 && kw > 0 && stride > 0, "bad arguments to p2pfl_im2col");
  const int Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
  P2PFL_REQUIRE(Ho > 0 && Wo > 0 && ld_col >= C * kh * kw, "bad im2col geometry");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  const int64_t total = (int64_t)B * Ho * Wo * C * kh * kw;
  im2col_kernel<<<grid_for(total, device), kThreads, 0, as_stream(stream)>>>(
      (const __nv_bfloat16*)x_nhwc, B, H, W, C, kh, kw, stride, pad, Ho, Wo, (__nv_bfloat16*)col, ld_col,
      (__nv_bfloat16*)col_t, ld_t);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_col2im(int device, void* stream, const void* dcol, int ld_col, int B, int H, int W, int C, int kh, int kw,
                 int stride, int pad, void* dx_nhwc) {
  P2PFL_REQUIRE(dcol && dx_nhwc, "bad arguments to p2pfl_col2im");
  const int Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
  P2PFL_CUDA_OK(cudaSetDevice(device));
  col2im_kernel<<<grid_for((int64_t)B * H * W * C, device), kThreads, 0, as_stream(stream)>>>(
      (const __nv_bfloat16*)dcol, ld_col, B, H, W, C, kh, kw, stride, pad, Ho, Wo, (__nv_bfloat16*)dx_nhwc);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_maxpool2_fwd(int device, void* stream, const void* x_nhwc, int B, int H, int W, int C, void* y_nhwc, void* arg,
                       void* flat, void* flat_t, int ld_flat_t) {
  P2PFL_REQUIRE(x_nhwc && arg && (H % 2) == 0 && (W % 2) == 0, "bad arguments to p2pfl_maxpool2_fwd");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  maxpool2_fwd_kernel<<<grid_for((int64_t)B * (H / 2) * (W / 2) * C, device), kThreads, 0, as_stream(stream)>>>(
      (const __nv_bfloat16*)x_nhwc, B, H, W, C, (__nv_bfloat16*)y_nhwc, (uint8_t*)arg, (__nv_bfloat16*)flat,
      (__nv_bfloat16*)flat_t, ld_flat_t);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_maxpool2_relu_bwd(int device, void* stream, const void* dy_nhwc, const void* dy_flat, const void* arg,
                            const void* act_nhwc, int B, int H, int W, int C, void* dx, void* dx_t, int64_t ld_t,
                            float* dbias) {
  P2PFL_REQUIRE((dy_nhwc || dy_flat) && arg && act_nhwc && dx, "bad arguments to p2pfl_maxpool2_relu_bwd");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  maxpool2_relu_bwd_kernel<<<grid_for((int64_t)B * H * W * C, device), kThreads, 0, as_stream(stream)>>>(
      (const __nv_bfloat16*)dy_nhwc, (const __nv_bfloat16*)dy_flat, (const uint8_t*)arg, (const __nv_bfloat16*)act_nhwc,
      B, H, W, C, (__nv_bfloat16*)dx, (__nv_bfloat16*)dx_t, ld_t, dbias);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_pad_cast(int device, void* stream, const float* src, int rows, int cols, void* dst_bf16, int ld) {
  P2PFL_REQUIRE(src && dst_bf16 && ld >= cols, "bad arguments to p2pfl_pad_cast");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  pad_cast_kernel<<<grid_for((int64_t)rows * ld, device), kThreads, 0, as_stream(stream)>>>(src, rows, cols,
                                                                                          (__nv_bfloat16*)dst_bf16, ld);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_nchw_to_nhwc(int device, void* stream, const float* x, int B, int C, int H, int W, void* y_bf16) {
  P2PFL_REQUIRE(x && y_bf16, "bad arguments to p2pfl_nchw_to_nhwc");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  nchw_to_nhwc_kernel<<<grid_for((int64_t)B * C * H * W, device), kThreads, 0, as_stream(stream)>>>(x, B, C, H, W,
                                                                                                 (__nv_bfloat16*)y_bf16);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_bn_fwd(int device, void* stream, const void* x_nhwc, int64_t pixels, int C, const float* gamma,
                 const float* beta, float* running_mean, float* running_var, float momentum, float eps, int relu,
                 const void* residual, void* y, float* save_mean, float* save_rstd) {
  P2PFL_REQUIRE(x_nhwc && y && gamma && beta && save_mean && save_rstd && pixels > 0 && C > 0, "bad arguments to p2pfl_bn_fwd");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  bn_fwd_kernel<<<C, kThreads, 0, as_stream(stream)>>>((const __nv_bfloat16*)x_nhwc, pixels, C, gamma, beta, running_mean,
                                                        running_var, momentum, eps, relu, (const __nv_bfloat16*)residual,
                                                        (__nv_bfloat16*)y, save_mean, save_rstd);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_bn_bwd(int device, void* stream, const void* x_nhwc, const void* y, const void* dy, int64_t pixels, int C,
                 const float* gamma, const float* save_mean, const float* save_rstd, int relu, void* dx,
                 void* d_residual, float* dgamma, float* dbeta) {
  P2PFL_REQUIRE(x_nhwc && dy && dx && gamma && save_mean && save_rstd && dgamma && dbeta, "bad arguments to p2pfl_bn_bwd");
  P2PFL_REQUIRE(!relu || y, "bn_bwd with relu needs the forward output");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  bn_bwd_kernel<<<C, kThreads, 0, as_stream(stream)>>>((const __nv_bfloat16*)x_nhwc, (const __nv_bfloat16*)y,
                                                        (const __nv_bfloat16*)dy, pixels, C, gamma, save_mean, save_rstd,
                                                        relu, (__nv_bfloat16*)dx, (__nv_bfloat16*)d_residual, dgamma, dbeta);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_avgpool_fwd(int device, void* stream, const void* x_nhwc, int B, int HW, int C, void* y, void* y_t, int ld_t) {
  P2PFL_REQUIRE(x_nhwc && y, "bad arguments to p2pfl_avgpool_fwd");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  avgpool_fwd_kernel<<<grid_for((int64_t)B * C, device), kThreads, 0, as_stream(stream)>>>(
      (const __nv_bfloat16*)x_nhwc, B, HW, C, (__nv_bfloat16*)y, (__nv_bfloat16*)y_t, ld_t);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_avgpool_bwd(int device, void* stream, const void* dy, int B, int HW, int C, void* dx_nhwc) {
  P2PFL_REQUIRE(dy && dx_nhwc, "bad arguments to p2pfl_avgpool_bwd");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  avgpool_bwd_kernel<<<grid_for((int64_t)B * HW * C, device), kThreads, 0, as_stream(stream)>>>(
      (const __nv_bfloat16*)dy, B, HW, C, (__nv_bfloat16*)dx_nhwc);
  P2PFL_LAUNCH_OK();
  return 0;
}

}  // extern "C"
