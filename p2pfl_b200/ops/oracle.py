"""ctypes binding of the plain C++ oracles (csrc/oracle/oracle.cpp -> lib/libp2pfl_oracle.so).

Each function is the specification of one kernel family in straight fp32 / fp64 loops; the GPU tests
compare the kernels with them (besides PyTorch) and ``bench/cpu_oracle_bench.py`` times them as the
"CPU doing the same work" baseline.  Inputs and outputs are contiguous fp32 / int32 torch CPU tensors.
"""

from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Tuple

import torch

_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib", "libp2pfl_oracle.so")
_lib: Optional[C.CDLL] = None


def available() -> bool:
    return os.path.exists(_PATH)


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not available():
            raise FileNotFoundError(f"{_PATH} is missing - build it with `python build.py`")
        _lib = C.CDLL(_PATH)
        _lib.oracle_softmax_xent.restype = C.c_double
        _lib.oracle_eval.restype = C.c_double
    return _lib


def _f(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to("cpu", torch.float32).contiguous()


def _p(t: Optional[torch.Tensor]) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def gemm_tn(A: torch.Tensor, B: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False) -> torch.Tensor:
    """D = A[M,K] . B[N,K]^T (+ bias per column) (ReLU)."""
    A, B = _f(A), _f(B)
    M, K = A.shape
    N = B.shape[0]
    D = torch.empty(M, N)
    b = None if bias is None else _f(bias)
    lib().oracle_gemm_tn(_p(A), _p(B), _p(D), M, N, K, K, K, N, _p(b), int(relu))
    return D


def softmax_xent(logits: torch.Tensor, labels: torch.Tensor) -> Tuple[float, torch.Tensor]:
    """(sum of per-sample losses, dlogits already divided by the batch size)."""
    x = _f(logits)
    y = labels.detach().to("cpu", torch.int32).contiguous()
    B, Cn = x.shape
    d = torch.empty(B, Cn)
    loss = lib().oracle_softmax_xent(_p(x), Cn, _p(y), B, Cn, _p(d), Cn)
    return float(loss), d


def adam(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, lr: float, t: int, beta1: float = 0.9,
         beta2: float = 0.999, eps: float = 1e-8) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:  # fmt: skip
    p, g, m, v = _f(p).clone(), _f(g), _f(m).clone(), _f(v).clone()
    adam_(p, g, m, v, lr, t, beta1, beta2, eps)
    return p, m, v


def adam_(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, lr: float, t: int, beta1: float = 0.9,
          beta2: float = 0.999, eps: float = 1e-8) -> None:  # fmt: skip
    """In place on contiguous fp32 CPU tensors (what the timed baseline calls: no copies around the loop)."""
    lib().oracle_adam(_p(p), _p(g), _p(m), _p(v), C.c_uint64(p.numel()), C.c_float(lr), C.c_float(beta1), C.c_float(beta2),
                      C.c_float(eps), C.c_int(t))  # fmt: skip


def fedavg(models: List[torch.Tensor], weights: List[float]) -> torch.Tensor:
    ms = [_f(x) for x in models]
    ptrs = (C.c_void_p * len(ms))(*[x.data_ptr() for x in ms])
    ws = (C.c_float * len(ms))(*weights)
    out = torch.empty_like(ms[0])
    lib().oracle_fedavg(ptrs, ws, len(ms), C.c_uint64(out.numel()), _p(out))
    return out


def conv2d(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], stride: int = 1, pad: int = 0) -> torch.Tensor:
    x, w = _f(x), _f(w)
    B, Cin, H, W = x.shape
    Cout, _, k, _ = w.shape
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    y = torch.empty(B, Cout, Ho, Wo)
    b = None if bias is None else _f(bias)
    lib().oracle_conv2d(_p(x), _p(w), _p(b), _p(y), B, Cin, H, W, Cout, k, stride, pad)
    return y


def maxpool2(x: torch.Tensor) -> torch.Tensor:
    x = _f(x)
    B, Cc, H, W = x.shape
    y = torch.empty(B, Cc, H // 2, W // 2)
    lib().oracle_maxpool2(_p(x), _p(y), B, Cc, H, W)
    return y


def batchnorm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = 1e-5, relu: bool = False) -> torch.Tensor:
    x = _f(x)
    B, Cc, H, W = x.shape
    y = torch.empty_like(x)
    lib().oracle_batchnorm(_p(x), _p(_f(gamma)), _p(_f(beta)), _p(y), B, Cc, H * W, C.c_float(eps), int(relu))
    return y


def evaluate(logits: torch.Tensor, labels: torch.Tensor) -> Tuple[float, int]:
    x = _f(logits)
    y = labels.detach().to("cpu", torch.int32).contiguous()
    B, Cn = x.shape
    correct = C.c_int32(0)
    loss = lib().oracle_eval(_p(x), Cn, _p(y), B, Cn, C.byref(correct))
    return float(loss), int(correct.value)
