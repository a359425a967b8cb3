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
//
// The ResNet plan uses the tap-major variants further down (im2col_kc / col2im_kc, K ordered
// (i, j, c), 16-byte channel vectors, weights repacked by conv_pack) and needs no transposed
// column matrix: its backward GEMMs read MN-major operands.
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
  }
}

// The transpose col_t[k, pix] is produced by its own pass with pix as the fastest index, so
// that the (much larger) writes are coalesced; the scattered side is the cached reads of x.
__global__ void __launch_bounds__(kThreads)
im2col_t_kernel(const __nv_bfloat16* __restrict__ x, int B, int H, int W, int C, int kh, int kw, int stride, int pad,
                int Ho, int Wo, __nv_bfloat16* __restrict__ col_t, int64_t ld_t) {
  const int K = C * kh * kw;
  const int64_t npix = (int64_t)B * Ho * Wo;
  const int64_t total = npix * K;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t pix = idx % npix;
    const int k = (int)(idx / npix);
    const int wo = (int)(pix % Wo);
    const int ho = (int)((pix / Wo) % Ho);
    const int b = (int)(pix / ((int64_t)Wo * Ho));
    const int c = k / (kh * kw);
    const int ij = k - c * kh * kw;
    const int i = ij / kw, j = ij - i * kw;
    const int h = ho * stride + i - pad, w = wo * stride + j - pad;
    __nv_bfloat16 v = __float2bfloat16_rn(0.f);
    if (h >= 0 && h < H && w >= 0 && w < W) v = x[(((int64_t)b * H + h) * W + w) * C + c];
    col_t[(int64_t)k * ld_t + pix] = v;
  }
}

// ------------------------------------------------------------------- col2im
// dx[b,h,w,c] = sum over (i,j) of dcol[(b,ho,wo), c*kh*kw + i*kw + j] with ho*s+i-p == h, wo*s+j-p == w.
// Gather form: one thread per input element, no atomics.
__global__ void __launch_bounds__(kThreads)
col2im_kernel(const __nv_bfloat16* __restrict__ dcol, int ld_col, int B, int H, int W, int C, int kh, int kw, int stride,
              int pad, int Ho, int Wo, __nv_bfloat16* __restrict__ dx, const __nv_bfloat16* __restrict__ extra) {
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
    if (extra) acc += __bfloat162float(extra[idx]);  // e.g. the residual branch's gradient
    dx[idx] = __float2bfloat16_rn(acc);
  }
}


// ----------------------------------------------- tap-major ("kc") conv layout
// For C % 8 == 0 the K axis of the implicit GEMM is ordered (i, j, c) - tap-major, channel
// fastest - instead of PyTorch's (c, i, j).  With NHWC activations every (pixel, tap) is then
// one contiguous run of C channels, so im2col / col2im move 16-byte vectors and the column
// matrix is written as one dense stream.  The weights are repacked to the same order once per
// optimizer step (conv_pack_kernel), the weight gradient is unpacked back (conv_unpack_kernel).
__global__ void __launch_bounds__(kThreads)
im2col_kc_kernel(const uint4* __restrict__ x, int B, int H, int W, int C8, int kh, int kw, int stride, int pad, int Ho,
                 int Wo, uint4* __restrict__ col, int ld8) {
  const int taps = kh * kw;
  const int64_t total = (int64_t)B * Ho * Wo * taps * C8;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int c8 = (int)(idx % C8);
    const int64_t r = idx / C8;
    const int tap = (int)(r % taps);
    const int64_t pix = r / taps;
    const int wo = (int)(pix % Wo);
    const int ho = (int)((pix / Wo) % Ho);
    const int b = (int)(pix / ((int64_t)Wo * Ho));
    const int i = tap / kw, j = tap - i * kw;
    const int h = ho * stride + i - pad, w = wo * stride + j - pad;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (h >= 0 && h < H && w >= 0 && w < W) v = x[(((int64_t)b * H + h) * W + w) * C8 + c8];
    col[pix * ld8 + (int64_t)tap * C8 + c8] = v;
  }
}

__device__ __forceinline__ void acc_bf16x8(float (&a)[8], const uint4& v) {
  const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    a[2 * k] += __uint_as_float(u[k] << 16);
    a[2 * k + 1] += __uint_as_float(u[k] & 0xFFFF0000u);
  }
}

__global__ void __launch_bounds__(kThreads)
col2im_kc_kernel(const uint4* __restrict__ dcol, int ld8, int B, int H, int W, int C8, int kh, int kw, int stride, int pad,
                 int Ho, int Wo, uint4* __restrict__ dx, const uint4* __restrict__ extra) {
  const int64_t total = (int64_t)B * H * W * C8;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int c8 = (int)(idx % C8);
    const int w = (int)((idx / C8) % W);
    const int h = (int)((idx / ((int64_t)C8 * W)) % H);
    const int b = (int)(idx / ((int64_t)C8 * W * H));
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
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
        acc_bf16x8(acc, dcol[pix * ld8 + (int64_t)(i * kw + j) * C8 + c8]);
      }
    }
    if (extra) acc_bf16x8(acc, extra[idx]);
    uint4 o;
    o.x = pack_bf16x2(acc[0], acc[1]);
    o.y = pack_bf16x2(acc[2], acc[3]);
    o.z = pack_bf16x2(acc[4], acc[5]);
    o.w = pack_bf16x2(acc[6], acc[7]);
    dx[idx] = o;
  }
}

// One launch for every convolution of a network; blockIdx.y picks the row of `table`:
//   {flat offset of the PyTorch weight [cout, cin*taps], offset in the packed buffer [cout, ldk],
//    cout, cin, taps, ldk, tap_major (1: k = tap*cin + c, 0: k = c*taps + tap), unused}
constexpr int kConvTableCols = 8;
// One thread per (cout, cin) pair walks the taps: its fp32 side is `taps` consecutive floats,
// its packed side is coalesced across the warp (consecutive c) - both directions stream.
__global__ void __launch_bounds__(kThreads)
conv_pack_kernel(const float* __restrict__ flat, __nv_bfloat16* __restrict__ packed, const int64_t* __restrict__ table) {
  const int64_t* t = table + (int64_t)blockIdx.y * kConvTableCols;
  const float* src = flat + t[0];
  __nv_bfloat16* dst = packed + t[1];
  const int cout = (int)t[2], cin = (int)t[3], taps = (int)t[4], ldk = (int)t[5], tap_major = (int)t[6];
  const int K = cin * taps;
  const int total = cout * cin;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int c = idx % cin;
    const int co = idx / cin;
    const float* s = src + (int64_t)co * K + c * taps;
    __nv_bfloat16* d = dst + (int64_t)co * ldk;
    for (int tap = 0; tap < taps; ++tap) d[tap_major ? tap * cin + c : c * taps + tap] = __float2bfloat16_rn(s[tap]);
  }
}

// gradient of the packed weight (fp32 [cout, ldk], what the wgrad GEMM accumulates) -> PyTorch order
__global__ void __launch_bounds__(kThreads)
conv_unpack_kernel(const float* __restrict__ packed, float* __restrict__ flat, const int64_t* __restrict__ table) {
  const int64_t* t = table + (int64_t)blockIdx.y * kConvTableCols;
  float* dst = flat + t[0];
  const float* src = packed + t[1];
  const int cout = (int)t[2], cin = (int)t[3], taps = (int)t[4], ldk = (int)t[5], tap_major = (int)t[6];
  const int K = cin * taps;
  const int total = cout * cin;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int c = idx % cin;
    const int co = idx / cin;
    float* d = dst + (int64_t)co * K + c * taps;
    const float* s = src + (int64_t)co * ldk;
    for (int tap = 0; tap < taps; ++tap) d[tap] = s[tap_major ? tap * cin + c : c * taps + tap];
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
// Training-mode BatchNorm2d over NHWC [pixels, C].  Every pass reads rows of C contiguous
// channels (coalesced): (1) per-channel sum / sum of squares with register partials and one
// atomic per channel per block, (2) a C-thread finalize (mean, rstd, running statistics),
// (3) an elementwise apply (+ residual, + ReLU).  Backward mirrors it: (1) per-channel
// sum(dy) and sum(dy * xhat) straight into dbeta / dgamma, (2) a tiled apply that writes dx
// in both orientations (the conv weight-gradient GEMM wants [C, pixels]).
__global__ void __launch_bounds__(kThreads)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, int64_t pixels, int C, float* __restrict__ sum,
                float* __restrict__ sumsq) {
  __shared__ float sh_s[kThreads], sh_q[kThreads];
  const int cpb = C < kThreads ? C : kThreads;  // channels per block pass
  const int rpp = kThreads / cpb;               // pixel rows per pass
  const int tx = threadIdx.x % cpb, ty = threadIdx.x / cpb;
  for (int cc = 0; cc < C; cc += cpb) {
    const int c = cc + tx;
    float s = 0.f, q = 0.f;
    if (c < C && ty < rpp) {
      for (int64_t p = (int64_t)blockIdx.x * rpp + ty; p < pixels; p += (int64_t)gridDim.x * rpp) {
        const float v = __bfloat162float(x[p * C + c]);
        s += v;
        q += v * v;
      }
    }
    sh_s[threadIdx.x] = s;
    sh_q[threadIdx.x] = q;
    __syncthreads();
    if (ty == 0 && c < C) {
      for (int r = 1; r < rpp; ++r) {
        s += sh_s[r * cpb + tx];
        q += sh_q[r * cpb + tx];
      }
      atomicAdd(sum + c, s);
      atomicAdd(sumsq + c, q);
    }
    __syncthreads();
  }
}

// sum / sumsq (in save_mean / save_rstd) -> mean / rstd in place, running statistics updated
__global__ void bn_finalize_kernel(float* __restrict__ save_mean, float* __restrict__ save_rstd, int64_t pixels, int C,
                                   float* __restrict__ running_mean, float* __restrict__ running_var, float momentum,
                                   float eps) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float mean = save_mean[c] / (float)pixels;
  const float var = fmaxf(save_rstd[c] / (float)pixels - mean * mean, 0.f);
  save_mean[c] = mean;
  save_rstd[c] = rsqrtf(var + eps);
  if (running_mean) {
    const float unbiased = pixels > 1 ? var * (float)pixels / (float)(pixels - 1) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
  }
}

// y = [relu]( (x - mean) * rstd * gamma + beta [+ residual] ); also used for inference with
// mean = running_mean, rstd computed from running_var by the caller-side kernel below
__global__ void __launch_bounds__(kThreads)
bn_apply_kernel(const __nv_bfloat16* __restrict__ x, int64_t total, int C, const float* __restrict__ gamma,
                const float* __restrict__ beta, const float* __restrict__ mean, const float* __restrict__ rstd_or_var,
                float eps, int var_given, int relu, const __nv_bfloat16* __restrict__ residual,
                __nv_bfloat16* __restrict__ y) {
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % C);
    const float rstd = var_given ? rsqrtf(rstd_or_var[c] + eps) : rstd_or_var[c];
    float v = (__bfloat162float(x[idx]) - mean[c]) * rstd * gamma[c] + beta[c];
    if (residual) v += __bfloat162float(residual[idx]);
    if (relu) v = fmaxf(v, 0.f);
    y[idx] = __float2bfloat16_rn(v);
  }
}

__global__ void __launch_bounds__(kThreads)
bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
                     const __nv_bfloat16* __restrict__ dy, int64_t pixels, int C, const float* __restrict__ save_mean,
                     const float* __restrict__ save_rstd, int relu, float* __restrict__ dgamma, float* __restrict__ dbeta) {
  __shared__ float sh_s[kThreads], sh_q[kThreads];
  const int cpb = C < kThreads ? C : kThreads;
  const int rpp = kThreads / cpb;
  const int tx = threadIdx.x % cpb, ty = threadIdx.x / cpb;
  for (int cc = 0; cc < C; cc += cpb) {
    const int c = cc + tx;
    float s = 0.f, q = 0.f;
    if (c < C && ty < rpp) {
      const float mean = save_mean[c], rstd = save_rstd[c];
      for (int64_t p = (int64_t)blockIdx.x * rpp + ty; p < pixels; p += (int64_t)gridDim.x * rpp) {
        float d = __bfloat162float(dy[p * C + c]);
        if (relu && !(__bfloat162float(y[p * C + c]) > 0.f)) d = 0.f;
        s += d;
        q += d * (__bfloat162float(x[p * C + c]) - mean) * rstd;
      }
    }
    sh_s[threadIdx.x] = s;
    sh_q[threadIdx.x] = q;
    __syncthreads();
    if (ty == 0 && c < C) {
      for (int r = 1; r < rpp; ++r) {
        s += sh_s[r * cpb + tx];
        q += sh_q[r * cpb + tx];
      }
      atomicAdd(dbeta + c, s);
      atomicAdd(dgamma + c, q);
    }
    __syncthreads();
  }
}

// 32 pixels x 32 channels per block (32 x 8 threads): dx row-major, and dx_t through a smem transpose
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
                    const __nv_bfloat16* __restrict__ dy, int64_t pixels, int C, const float* __restrict__ gamma,
                    const float* __restrict__ save_mean, const float* __restrict__ save_rstd, int relu,
                    const float* __restrict__ dgamma, const float* __restrict__ dbeta, __nv_bfloat16* __restrict__ dx,
                    __nv_bfloat16* __restrict__ dx_t, int64_t ld_t, __nv_bfloat16* __restrict__ d_residual) {
  __shared__ float tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int tiles_c = (C + 31) / 32;
  const int64_t tiles_p = (pixels + 31) / 32;
  const float inv_n = 1.f / (float)pixels;
  for (int64_t t = blockIdx.x; t < tiles_p * tiles_c; t += gridDim.x) {
    const int64_t p0 = (t / tiles_c) * 32;
    const int c = (int)(t % tiles_c) * 32 + tx;
    float g = 0.f, mean = 0.f, rstd = 0.f, sdy = 0.f, sdyx = 0.f;
    if (c < C) {
      g = gamma[c];
      mean = save_mean[c];
      rstd = save_rstd[c];
      sdy = dbeta[c];
      sdyx = dgamma[c];
    }
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      const int64_t p = p0 + ty + j;
      float out = 0.f;
      if (p < pixels && c < C) {
        const int64_t idx = p * C + c;
        float d = __bfloat162float(dy[idx]);
        if (relu && !(__bfloat162float(y[idx]) > 0.f)) d = 0.f;
        if (d_residual) d_residual[idx] = __float2bfloat16_rn(d);
        const float xh = (__bfloat162float(x[idx]) - mean) * rstd;
        out = g * rstd * (d - sdy * inv_n - xh * sdyx * inv_n);
        dx[idx] = __float2bfloat16_rn(out);
      }
      tile[ty + j][tx] = out;
    }
    if (dx_t) {
      __syncthreads();
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        const int cc = (int)(t % tiles_c) * 32 + ty + j;
        const int64_t p = p0 + tx;
        if (cc < C && p < pixels) dx_t[(int64_t)cc * ld_t + p] = __float2bfloat16_rn(tile[tx][ty + j]);
      }
    }
    __syncthreads();
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
      (__nv_bfloat16*)nullptr, ld_t);
  P2PFL_LAUNCH_OK();
  if (col_t) {
    im2col_t_kernel<<<grid_for(total, device), kThreads, 0, as_stream(stream)>>>(
        (const __nv_bfloat16*)x_nhwc, B, H, W, C, kh, kw, stride, pad, Ho, Wo, (__nv_bfloat16*)col_t, ld_t);
    P2PFL_LAUNCH_OK();
  }
  return 0;
}

int p2pfl_col2im(int device, void* stream, const void* dcol, int ld_col, int B, int H, int W, int C, int kh, int kw,
                 int stride, int pad, void* dx_nhwc) {
  return p2pfl_col2im_add(device, stream, dcol, ld_col, B, H, W, C, kh, kw, stride, pad, dx_nhwc, nullptr);
}

int p2pfl_col2im_add(int device, void* stream, const void* dcol, int ld_col, int B, int H, int W, int C, int kh, int kw,
                     int stride, int pad, void* dx_nhwc, const void* extra) {
  P2PFL_REQUIRE(dcol && dx_nhwc, "bad arguments to p2pfl_col2im");
  const int Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
  P2PFL_CUDA_OK(cudaSetDevice(device));
  col2im_kernel<<<grid_for((int64_t)B * H * W * C, device), kThreads, 0, as_stream(stream)>>>(
      (const __nv_bfloat16*)dcol, ld_col, B, H, W, C, kh, kw, stride, pad, Ho, Wo, (__nv_bfloat16*)dx_nhwc,
      (const __nv_bfloat16*)extra);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_im2col_kc(int device, void* stream, const void* x_nhwc, int B, int H, int W, int C, int kh, int kw, int stride,
                    int pad, void* col, int ld_col) {
  P2PFL_REQUIRE(x_nhwc && col && C > 0 && (C % 8) == 0 && (ld_col % 8) == 0 && ld_col >= C * kh * kw,
                "p2pfl_im2col_kc needs C and ld_col to be multiples of 8");
  P2PFL_REQUIRE(((reinterpret_cast<uintptr_t>(x_nhwc) | reinterpret_cast<uintptr_t>(col)) & 15) == 0,
                "p2pfl_im2col_kc needs 16-byte aligned buffers");
  const int Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
  P2PFL_CUDA_OK(cudaSetDevice(device));
  const int64_t total = (int64_t)B * Ho * Wo * kh * kw * (C / 8);
  im2col_kc_kernel<<<grid_for(total, device), kThreads, 0, as_stream(stream)>>>(
      (const uint4*)x_nhwc, B, H, W, C / 8, kh, kw, stride, pad, Ho, Wo, (uint4*)col, ld_col / 8);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_col2im_kc(int device, void* stream, const void* dcol, int ld_col, int B, int H, int W, int C, int kh, int kw,
                    int stride, int pad, void* dx_nhwc, const void* extra) {
  P2PFL_REQUIRE(dcol && dx_nhwc && C > 0 && (C % 8) == 0 && (ld_col % 8) == 0, "p2pfl_col2im_kc needs C and ld_col to be multiples of 8");
  P2PFL_REQUIRE(((reinterpret_cast<uintptr_t>(dcol) | reinterpret_cast<uintptr_t>(dx_nhwc) | reinterpret_cast<uintptr_t>(extra)) & 15) == 0,
                "p2pfl_col2im_kc needs 16-byte aligned buffers");
  const int Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
  P2PFL_CUDA_OK(cudaSetDevice(device));
  col2im_kc_kernel<<<grid_for((int64_t)B * H * W * (C / 8), device), kThreads, 0, as_stream(stream)>>>(
      (const uint4*)dcol, ld_col / 8, B, H, W, C / 8, kh, kw, stride, pad, Ho, Wo, (uint4*)dx_nhwc, (const uint4*)extra);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_conv_pack(int device, void* stream, const float* flat, void* packed_bf16, const int64_t* table_dev, int n_conv,
                    int64_t max_elems) {
  P2PFL_REQUIRE(flat && packed_bf16 && table_dev && n_conv > 0 && n_conv <= 65535, "bad arguments to p2pfl_conv_pack");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  int gx = grid_for(max_elems / 9 + 1, device);
  gx = gx > 256 ? 256 : gx;
  conv_pack_kernel<<<dim3(gx, n_conv), kThreads, 0, as_stream(stream)>>>(flat, (__nv_bfloat16*)packed_bf16, table_dev);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_conv_unpack(int device, void* stream, const float* packed_f32, float* flat, const int64_t* table_dev, int n_conv,
                      int64_t max_elems) {
  P2PFL_REQUIRE(flat && packed_f32 && table_dev && n_conv > 0 && n_conv <= 65535, "bad arguments to p2pfl_conv_unpack");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  int gx = grid_for(max_elems / 9 + 1, device);
  gx = gx > 256 ? 256 : gx;
  conv_unpack_kernel<<<dim3(gx, n_conv), kThreads, 0, as_stream(stream)>>>(packed_f32, flat, table_dev);
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
  cudaStream_t s = as_stream(stream);
  P2PFL_CUDA_OK(cudaMemsetAsync(save_mean, 0, sizeof(float) * C, s));
  P2PFL_CUDA_OK(cudaMemsetAsync(save_rstd, 0, sizeof(float) * C, s));
  const int rpp = kThreads / (C < kThreads ? C : kThreads);
  int64_t blocks = (pixels + rpp - 1) / rpp / 8 + 1;
  const int64_t cap = (int64_t)sm_count(device) * 4;
  bn_stats_kernel<<<(int)(blocks < cap ? blocks : cap), kThreads, 0, s>>>((const __nv_bfloat16*)x_nhwc, pixels, C, save_mean,
                                                                           save_rstd);
  P2PFL_LAUNCH_OK();
  bn_finalize_kernel<<<(C + 127) / 128, 128, 0, s>>>(save_mean, save_rstd, pixels, C, running_mean, running_var, momentum, eps);
  P2PFL_LAUNCH_OK();
  bn_apply_kernel<<<grid_for(pixels * C, device), kThreads, 0, s>>>((const __nv_bfloat16*)x_nhwc, pixels * C, C, gamma, beta,
                                                                     save_mean, save_rstd, eps, 0, relu,
                                                                     (const __nv_bfloat16*)residual, (__nv_bfloat16*)y);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_bn_infer(int device, void* stream, const void* x_nhwc, int64_t pixels, int C, const float* gamma,
                   const float* beta, const float* running_mean, const float* running_var, float eps, int relu,
                   const void* residual, void* y) {
  P2PFL_REQUIRE(x_nhwc && y && gamma && beta && running_mean && running_var, "bad arguments to p2pfl_bn_infer");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  bn_apply_kernel<<<grid_for(pixels * C, device), kThreads, 0, as_stream(stream)>>>(
      (const __nv_bfloat16*)x_nhwc, pixels * C, C, gamma, beta, running_mean, running_var, eps, 1, relu,
      (const __nv_bfloat16*)residual, (__nv_bfloat16*)y);
  P2PFL_LAUNCH_OK();
  return 0;
}

int p2pfl_bn_bwd(int device, void* stream, const void* x_nhwc, const void* y, const void* dy, int64_t pixels, int C,
                 const float* gamma, const float* save_mean, const float* save_rstd, int relu, void* dx,
                 void* d_residual, float* dgamma, float* dbeta) {
  return p2pfl_bn_bwd_t(device, stream, x_nhwc, y, dy, pixels, C, gamma, save_mean, save_rstd, relu, dx, nullptr, 0,
                        d_residual, dgamma, dbeta);
}

int p2pfl_bn_bwd_t(int device, void* stream, const void* x_nhwc, const void* y, const void* dy, int64_t pixels, int C,
                   const float* gamma, const float* save_mean, const float* save_rstd, int relu, void* dx, void* dx_t,
                   int64_t ld_t, void* d_residual, float* dgamma, float* dbeta) {
  P2PFL_REQUIRE(x_nhwc && dy && dx && gamma && save_mean && save_rstd && dgamma && dbeta, "bad arguments to p2pfl_bn_bwd");
  P2PFL_REQUIRE(!relu || y, "bn_bwd with relu needs the forward output");
  P2PFL_CUDA_OK(cudaSetDevice(device));
  cudaStream_t s = as_stream(stream);
  P2PFL_CUDA_OK(cudaMemsetAsync(dgamma, 0, sizeof(float) * C, s));
  P2PFL_CUDA_OK(cudaMemsetAsync(dbeta, 0, sizeof(float) * C, s));
  const int rpp = kThreads / (C < kThreads ? C : kThreads);
  int64_t blocks = (pixels + rpp - 1) / rpp / 8 + 1;
  const int64_t cap = (int64_t)sm_count(device) * 4;
  bn_bwd_reduce_kernel<<<(int)(blocks < cap ? blocks : cap), kThreads, 0, s>>>(
      (const __nv_bfloat16*)x_nhwc, (const __nv_bfloat16*)y, (const __nv_bfloat16*)dy, pixels, C, save_mean, save_rstd, relu,
      dgamma, dbeta);
  P2PFL_LAUNCH_OK();
  const int64_t tiles = ((pixels + 31) / 32) * ((C + 31) / 32);
  const int64_t cap2 = (int64_t)sm_count(device) * 8;
  bn_bwd_apply_kernel<<<(int)(tiles < cap2 ? tiles : cap2), 256, 0, s>>>(
      (const __nv_bfloat16*)x_nhwc, (const __nv_bfloat16*)y, (const __nv_bfloat16*)dy, pixels, C, gamma, save_mean, save_rstd,
      relu, dgamma, dbeta, (__nv_bfloat16*)dx, (__nv_bfloat16*)dx_t, ld_t, (__nv_bfloat16*)d_residual);
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
