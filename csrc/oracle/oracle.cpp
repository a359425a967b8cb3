// Plain C++ oracles and timed CPU baselines of the operations the sm_100a kernels implement
// (one oracle per kernel row).  Straight loops in fp32 (fp64
// accumulation where a kernel accumulates in fp32 over long reductions), no vectorisation tricks:
// they are the specification the kernels are compared with, and - timed - the "what would the CPU
// do" row next to each kernel's number.  Built by build.py with g++ into libp2pfl_oracle.so.
#include <cmath>
#include <cstdint>
#include <cstring>

extern "C" {

// D[M,N] = A[M,K] . B[N,K]^T (+ bias per column) (ReLU)      reference: mlp.py:63-67 (nn.Linear + relu)
void oracle_gemm_tn(const float* A, const float* B, float* D, int M, int N, int K, int lda, int ldb, int ldd,
                    const float* bias, int relu) {
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      double acc = 0.0;
      for (int k = 0; k < K; ++k) acc += (double)A[(size_t)m * lda + k] * (double)B[(size_t)n * ldb + k];
      float v = (float)acc + (bias ? bias[n] : 0.f);
      D[(size_t)m * ldd + n] = (relu && v < 0.f) ? 0.f : v;
    }
}

// log-softmax + mean cross-entropy forward and backward      reference: mlp.py:68,78 (log_softmax + cross_entropy)
// dlogits = (softmax - onehot) / B; returns the SUM of the per-sample losses
double oracle_softmax_xent(const float* logits, int ld, const int32_t* labels, int B, int C, float* dlogits, int ld_d) {
  double total = 0.0;
  for (int b = 0; b < B; ++b) {
    const float* x = logits + (size_t)b * ld;
    float mx = x[0];
    for (int c = 1; c < C; ++c) mx = x[c] > mx ? x[c] : mx;
    double se = 0.0;
    for (int c = 0; c < C; ++c) se += std::exp((double)x[c] - mx);
    const double lse = mx + std::log(se);
    total += lse - x[labels[b]];
    if (dlogits)
      for (int c = 0; c < C; ++c)
        dlogits[(size_t)b * ld_d + c] = (float)((std::exp((double)x[c] - lse) - (c == labels[b] ? 1.0 : 0.0)) / B);
  }
  return total;
}

// torch.optim.Adam.step, weight decay 0                      reference: mlp.py:71-73
void oracle_adam(float* p, const float* g, float* m, float* v, uint64_t n, float lr, float b1, float b2, float eps, int t) {
  const double bc1 = 1.0 - std::pow((double)b1, t), bc2 = 1.0 - std::pow((double)b2, t);
  for (uint64_t i = 0; i < n; ++i) {
    m[i] = b1 * m[i] + (1.f - b1) * g[i];
    v[i] = b2 * v[i] + (1.f - b2) * g[i] * g[i];
    const double denom = std::sqrt((double)v[i]) / std::sqrt(bc2) + eps;
    p[i] = (float)(p[i] - (lr / bc1) * (m[i] / denom));
  }
}

// out = sum_i w[i] * model_i (weights already normalised)    reference: fedavg.py:49-58
void oracle_fedavg(const float* const* models, const float* w, int n_models, uint64_t n, float* out) {
  for (uint64_t i = 0; i < n; ++i) {
    double acc = 0.0;
    for (int k = 0; k < n_models; ++k) acc += (double)w[k] * models[k][i];
    out[i] = (float)acc;
  }
}

// direct convolution, NCHW, square kernel, zero padding      reference: cnn.py:78-81 (nn.Conv2d)
void oracle_conv2d(const float* x, const float* w, const float* bias, float* y, int B, int Cin, int H, int W, int Cout,
                   int k, int stride, int pad) {
  const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  for (int b = 0; b < B; ++b)
    for (int co = 0; co < Cout; ++co)
      for (int ho = 0; ho < Ho; ++ho)
        for (int wo = 0; wo < Wo; ++wo) {
          double acc = bias ? bias[co] : 0.0;
          for (int ci = 0; ci < Cin; ++ci)
            for (int i = 0; i < k; ++i)
              for (int j = 0; j < k; ++j) {
                const int h = ho * stride + i - pad, ww = wo * stride + j - pad;
                if (h < 0 || h >= H || ww < 0 || ww >= W) continue;
                acc += (double)x[(((size_t)b * Cin + ci) * H + h) * W + ww] * w[(((size_t)co * Cin + ci) * k + i) * k + j];
              }
          y[(((size_t)b * Cout + co) * Ho + ho) * Wo + wo] = (float)acc;
        }
}

// 2x2 / stride 2 max pooling, NCHW                           reference: cnn.py:79,81 (nn.MaxPool2d(2))
void oracle_maxpool2(const float* x, float* y, int B, int C, int H, int W) {
  const int Ho = H / 2, Wo = W / 2;
  for (int bc = 0; bc < B * C; ++bc)
    for (int ho = 0; ho < Ho; ++ho)
      for (int wo = 0; wo < Wo; ++wo) {
        const float* p = x + ((size_t)bc * H + 2 * ho) * W + 2 * wo;
        float m = p[0];
        m = p[1] > m ? p[1] : m;
        m = p[W] > m ? p[W] : m;
        m = p[W + 1] > m ? p[W + 1] : m;
        y[((size_t)bc * Ho + ho) * Wo + wo] = m;
      }
}

// training-mode BatchNorm2d forward over NCHW (biased variance), optional ReLU     (ResNet-18 of config 3)
void oracle_batchnorm(const float* x, const float* gamma, const float* beta, float* y, int B, int C, int HW, float eps, int relu) {
  for (int c = 0; c < C; ++c) {
    double s = 0.0, ss = 0.0;
    for (int b = 0; b < B; ++b)
      for (int i = 0; i < HW; ++i) {
        const double v = x[((size_t)b * C + c) * HW + i];
        s += v;
        ss += v * v;
      }
    const double n = (double)B * HW, mean = s / n, var = ss / n - mean * mean, rstd = 1.0 / std::sqrt(var + eps);
    for (int b = 0; b < B; ++b)
      for (int i = 0; i < HW; ++i) {
        const size_t idx = ((size_t)b * C + c) * HW + i;
        float v = (float)((x[idx] - mean) * rstd * gamma[c] + beta[c]);
        y[idx] = (relu && v < 0.f) ? 0.f : v;
      }
  }
}

// evaluation reduce: sum of per-sample losses and number of correct argmax predictions   reference: mlp.py:96-99
double oracle_eval(const float* logits, int ld, const int32_t* labels, int B, int C, int32_t* correct) {
  int ok = 0;
  for (int b = 0; b < B; ++b) {
    const float* x = logits + (size_t)b * ld;
    int arg = 0;
    for (int c = 1; c < C; ++c) arg = x[c] > x[arg] ? c : arg;
    ok += arg == labels[b];
  }
  if (correct) *correct = ok;
  return oracle_softmax_xent(logits, ld, labels, B, C, nullptr, 0);
}

}  // extern "C"
