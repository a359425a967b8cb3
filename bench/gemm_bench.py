#!/usr/bin/env python
"""tcgen05 GEMM (ops.gemm_tn) vs cuBLAS (torch.matmul, bf16) on the shapes the
learners use and on large square shapes.  CUDA-event timing, L2 flushed between
iterations, best/mean of 20.  One JSON line per shape."""

import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from p2pfl_b200 import ops  # noqa: E402

SHAPES = [
    ("mlp.l1 fwd", 256, 32, 784), ("mlp.l2 fwd", 128, 32, 256), ("mlp.l3 fwd", 10, 32, 128),
    ("mlp.l1 wgrad", 256, 784, 32), ("mlp.l2 dgrad", 256, 32, 128),
    ("cnn.fc1 fwd", 2048, 32, 3136), ("cnn.fc1 wgrad", 2048, 3136, 32),
    ("eval l1 (N=1250)", 256, 1250, 784),
    ("square 2048", 2048, 2048, 2048), ("square 4096", 4096, 4096, 4096), ("square 8192", 8192, 8192, 8192),
]  # fmt: skip


def timeit(fn, flush, iters=20):
    ts = []
    for _ in range(3):
        fn()
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts), sum(ts) / len(ts)


def main():
    dev = "cuda:0"
    flush = torch.empty(64 << 20, dtype=torch.float32, device=dev)
    for name, M, N, K in SHAPES:
        ld = (K + 7) // 8 * 8
        A = torch.zeros(M, ld, device=dev, dtype=torch.bfloat16)
        B = torch.zeros(N, ld, device=dev, dtype=torch.bfloat16)
        A[:, :K] = (torch.randn(M, K, device=dev) / K**0.5).to(torch.bfloat16)
        B[:, :K] = torch.randn(N, K, device=dev).to(torch.bfloat16)
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        mine = timeit(lambda: ops.gemm_tn(A, ld, B, ld, M, N, K, out_bf16=out, ld_bf16=N), flush)
        Ac, Bc = A[:, :K].contiguous(), B[:, :K].contiguous()
        ref = timeit(lambda: torch.matmul(Ac, Bc.t()), flush)
        err = (out.float() - (Ac.float() @ Bc.float().t())).abs().max().item()
        flops = 2.0 * M * N * K
        print(json.dumps({
            "shape": name, "M": M, "N": N, "K": K, "tcgen05_us_best": mine[0] * 1e3, "tcgen05_us_mean": mine[1] * 1e3,
            "cublas_us_best": ref[0] * 1e3, "cublas_us_mean": ref[1] * 1e3,
            "tcgen05_tflops": flops / (mine[0] * 1e-3) / 1e12, "cublas_tflops": flops / (ref[0] * 1e-3) / 1e12,
            "max_abs_err": err,
        }), flush=True)


if __name__ == "__main__":
    main()
