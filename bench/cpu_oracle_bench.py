#!/usr/bin/env python
"""Timed CPU baseline: the C++ oracles (csrc/oracle, one thread, -O2) doing the work of one MLP train
step, one FedAvg of 8 models and one Adam step - the rows next to the kernels' numbers in BASELINE.md.
No GPU needed.  Prints one JSON line."""
import json, os, sys, time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import build  # noqa: E402

build.build_oracle()
from p2pfl_b200.ops import oracle as o  # noqa: E402


def best(fn, n=5):
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return min(ts) * 1e3


torch.set_num_threads(1)
torch.manual_seed(0)
B = 32
X, W1, W2, W3 = torch.randn(B, 784), torch.randn(256, 784), torch.randn(128, 256), torch.randn(10, 128)
y = torch.randint(0, 10, (B,))


def mlp_step():
    h1 = o.gemm_tn(X, W1, relu=True)
    h2 = o.gemm_tn(h1, W2, relu=True)
    z = o.gemm_tn(h2, W3)
    _, dz = o.softmax_xent(z, y)
    dh2 = o.gemm_tn(dz, W3.t().contiguous())
    dh1 = o.gemm_tn(dh2, W2.t().contiguous())
    o.gemm_tn(dz.t().contiguous(), h2.t().contiguous())      # dW3
    o.gemm_tn(dh2.t().contiguous(), h1.t().contiguous())     # dW2
    o.gemm_tn(dh1.t().contiguous(), X.t().contiguous())      # dW1


n = 235168
p, g, m, v = (torch.randn(n) for _ in range(4))
v = v.abs()
models = [torch.randn(n) for _ in range(8)]
out = {
    "bench": "cpu_oracle", "threads": 1,
    "mlp_step_fwd_bwd_ms": best(mlp_step, 3),
    "adam_235k_ms": best(lambda: o.adam_(p, g, m, v, 1e-3, 3)),
    "fedavg_8x235k_ms": best(lambda: o.fedavg(models, [0.125] * 8)),
    "note": "C++ -O2 straight loops (the oracles the kernels are tested against); the epoch kernel does step + Adam in 0.0117 ms, "
            "the fused exchange does push + FedAvg of 8 models in 0.075 ms",
}
print(json.dumps(out), flush=True)
