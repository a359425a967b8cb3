"""Python bindings of the sm_100a kernel library (``lib/libp2pfl_b200.so``).

The library exposes a plain C ABI (``include/p2pfl_b200.h``) and is loaded with
``ctypes``; tensors are passed as raw device pointers and the launch goes on
the *current torch stream*, so kernels compose with torch ops, CUDA events and
CUDA-graph capture.  There is no fallback: on a machine with a GPU a missing
library is an error, never a silent eager path.

``launch_count()`` reports how many of this library's kernels were launched
from Python (graph replays add their node count) - bench.py's ``gpu_launches``.
"""

from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Any, List, Optional, Sequence, Tuple

import torch

from p2pfl_b200.exceptions import ExtensionMissingError

_LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib", "libp2pfl_b200.so")
_lib: Optional[C.CDLL] = None
_lock = threading.Lock()
_launches = 0

MAX_PEERS = 16
MAX_MODELS = 16


# --------------------------------------------------------------------- structs
class GemmArgs(C.Structure):
    _fields_ = [
        ("A", C.c_void_p), ("lda", C.c_int32),
        ("B", C.c_void_p), ("ldb", C.c_int32),
        ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32),
        ("bias", C.c_void_p), ("bias_per_row", C.c_int32),
        ("relu", C.c_int32),
        ("mask", C.c_void_p), ("ld_mask", C.c_int32),
        ("out_f32", C.c_void_p), ("ld_f32", C.c_int32),
        ("out_bf16", C.c_void_p), ("ld_bf16", C.c_int32),
        ("out_f32_t", C.c_void_p), ("ld_f32_t", C.c_int32),
        ("out_bf16_t", C.c_void_p), ("ld_bf16_t", C.c_int32),
        ("rowsum", C.c_void_p),
        ("alpha", C.c_float),
        ("split_k", C.c_int32), ("mn_major", C.c_int32),
    ]  # fmt: skip


class CeArgs(C.Structure):
    _fields_ = [
        ("logits", C.c_void_p), ("ld_logits", C.c_int32),
        ("labels", C.c_void_p), ("B", C.c_int32), ("C", C.c_int32),
        ("dlogits", C.c_void_p), ("ld_d", C.c_int32),
        ("dlogits_t", C.c_void_p), ("ld_dt", C.c_int32),
        ("dbias", C.c_void_p),
        ("loss_sum", C.c_void_p),
    ]  # fmt: skip


class PushTable(C.Structure):
    _fields_ = [
        ("inbox", C.c_void_p * MAX_PEERS),
        ("flags", C.c_void_p * MAX_PEERS),
        ("n_peers", C.c_int32),
        ("wire_bf16", C.c_int32),
    ]


class ReduceTable(C.Structure):
    _fields_ = [
        ("slot", C.c_void_p * MAX_PEERS),
        ("flags", C.c_void_p * MAX_PEERS),
        ("weight", C.c_float * MAX_PEERS),
        ("n_slots", C.c_int32),
        ("self_pos", C.c_int32),
        ("self_weight", C.c_float),
        ("wire_bf16", C.c_int32),
        ("round_self", C.c_int32),
        ("pad_", C.c_int32),
    ]


class TransposeDesc(C.Structure):
    _fields_ = [
        ("src_off", C.c_uint64),
        ("dst_off", C.c_uint64),
        ("rows", C.c_int32),
        ("cols", C.c_int32),
        ("ld_t", C.c_int32),
        ("pad_", C.c_int32),
    ]


# ---------------------------------------------------------------------- loader
def library_path() -> str:
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(_LIB_PATH):
                    raise ExtensionMissingError(
                        f"{_LIB_PATH} is missing - build it with `python build.py` (see __graft_entry__.build)"
                    )
                l = C.CDLL(_LIB_PATH)
                l.p2pfl_last_error.restype = C.c_char_p
                _lib = l
    return _lib


def _check(rc: int) -> None:
    if rc != 0:
        msg = lib().p2pfl_last_error().decode(errors="replace")
        raise RuntimeError(f"p2pfl_b200 kernel library error {rc}: {msg}")


def _stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dev(t: torch.Tensor) -> int:
    return t.device.index if t.device.index is not None else torch.cuda.current_device()


def _p(t: Optional[torch.Tensor]) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def _count(n: int = 1) -> None:
    global _launches
    _launches += n


def set_pdl(enabled: bool) -> bool:
    """Programmatic dependent launch for the train-step kernels; returns the previous setting."""
    return bool(lib().p2pfl_set_pdl(C.c_int(int(enabled))))


def launch_count() -> int:
    return _launches


def add_launches(n: int) -> None:
    """Account for kernels replayed from a CUDA graph."""
    _count(n)


# ------------------------------------------------------------------ optimizers
def adam_step(
    p: torch.Tensor,
    g: torch.Tensor,
    m: torch.Tensor,
    v: torch.Tensor,
    p_bf16: Optional[torch.Tensor],
    lr: float,
    step_ptr: Optional[torch.Tensor],
    step_offset: int,
    beta1: float = 0.9,
    beta2: float = 0.999,
    eps: float = 1e-8,
    weight_decay: float = 0.0,
) -> None:
    _check(
        lib().p2pfl_adam_step(
            _dev(p), _stream(), _p(p), _p(g), _p(m), _p(v), _p(p_bf16), C.c_uint64(p.numel()),
            C.c_float(lr), C.c_float(beta1), C.c_float(beta2), C.c_float(eps), C.c_float(weight_decay),
            _p(step_ptr), C.c_int32(step_offset),
        )
    )  # fmt: skip
    _count()


def adam_step_t(
    p, g, m, v, p_bf16, lr: float, step_ptr, step_offset: int, wt_bf16, descs: Sequence[tuple],
    beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8, weight_decay: float = 0.0,
) -> None:  # fmt: skip
    """Adam + bf16 refresh + bf16 transposed weight copies (descs as in transpose_cast) in one launch."""
    arr = (TransposeDesc * max(1, len(descs)))()
    for j, (so, do, r, c, ld) in enumerate(descs):
        arr[j] = TransposeDesc(so, do, r, c, ld, 0)
    _check(
        lib().p2pfl_adam_step_t(
            _dev(p), _stream(), _p(p), _p(g), _p(m), _p(v), _p(p_bf16), C.c_uint64(p.numel()),
            C.c_float(lr), C.c_float(beta1), C.c_float(beta2), C.c_float(eps), C.c_float(weight_decay),
            _p(step_ptr), C.c_int32(step_offset), _p(wt_bf16), arr, C.c_int(len(descs)),
        )
    )  # fmt: skip
    _count()


def sgd_step(p, g, mom, p_bf16, lr: float, momentum: float = 0.0) -> None:
    _check(
        lib().p2pfl_sgd_step(
            _dev(p), _stream(), _p(p), _p(g), _p(mom), _p(p_bf16), C.c_uint64(p.numel()), C.c_float(lr), C.c_float(momentum)
        )
    )
    _count()


def add_i32(t: torch.Tensor, delta: int) -> None:
    _check(lib().p2pfl_add_i32(_dev(t), _stream(), _p(t), C.c_int32(delta)))
    _count()


def cast_f32_bf16(src: torch.Tensor, dst: torch.Tensor) -> None:
    _check(lib().p2pfl_cast_f32_bf16(_dev(src), _stream(), _p(src), _p(dst), C.c_uint64(src.numel())))
    _count()


def transpose_cast(src: torch.Tensor, dst: torch.Tensor, descs: Sequence[tuple]) -> None:
    """descs: (src_off, dst_off, rows, cols, ld_t) per matrix, at most 16 per call."""
    for i in range(0, len(descs), 16):
        chunk = descs[i : i + 16]
        arr = (TransposeDesc * len(chunk))()
        for j, (so, do, r, c, ld) in enumerate(chunk):
            arr[j] = TransposeDesc(so, do, r, c, ld, 0)
        _check(lib().p2pfl_transpose_cast(_dev(src), _stream(), _p(src), _p(dst), arr, C.c_int(len(chunk))))
        _count()


# ---------------------------------------------------------------------- losses
def softmax_xent(
    logits: torch.Tensor,
    ld_logits: int,
    labels: torch.Tensor,
    B: int,
    Cn: int,
    dlogits: Optional[torch.Tensor],
    ld_d: int,
    dlogits_t: Optional[torch.Tensor],
    ld_dt: int,
    dbias: Optional[torch.Tensor],
    loss_sum: Optional[torch.Tensor],
) -> None:
    _check(
        lib().p2pfl_softmax_xent(
            _dev(logits), _stream(), _p(logits), C.c_int(ld_logits), _p(labels), C.c_int(B), C.c_int(Cn),
            _p(dlogits), C.c_int(ld_d), _p(dlogits_t), C.c_int(ld_dt), _p(dbias), _p(loss_sum),
        )
    )  # fmt: skip
    _count()


def eval_reduce(logits: torch.Tensor, ld_logits: int, labels: torch.Tensor, B: int, Cn: int, loss_sum, correct) -> None:
    _check(
        lib().p2pfl_eval_reduce(
            _dev(logits), _stream(), _p(logits), C.c_int(ld_logits), _p(labels), C.c_int(B), C.c_int(Cn), _p(loss_sum), _p(correct)
        )
    )
    _count()


# ---------------------------------------------------------------------- FedAvg
def fedavg_into(
    models: List[torch.Tensor], weights: List[float], out_f32: Optional[torch.Tensor], out_bf16: Optional[torch.Tensor]
) -> None:
    """out = sum_i weights[i] * models[i]; at most 16 models per launch (longer
    lists are folded: the partial result re-enters with weight 1)."""
    n = models[0].numel()
    dev = _dev(models[0])
    models, weights = list(models), list(weights)
    while models:
        k = min(len(models), MAX_MODELS)
        last = k == len(models)
        ptrs = (C.c_void_p * k)(*[m.data_ptr() for m in models[:k]])
        ws = (C.c_float * k)(*weights[:k])
        isb = (C.c_int32 * k)(*[1 if m.dtype == torch.bfloat16 else 0 for m in models[:k]])
        tmp = out_f32 if (out_f32 is not None) else torch.empty(n, dtype=torch.float32, device=models[0].device)
        _check(
            lib().p2pfl_fedavg(dev, _stream(), C.c_int(k), ptrs, ws, isb, _p(tmp), _p(out_bf16 if last else None), C.c_uint64(n))
        )
        _count()
        if last:
            break
        models = [tmp] + models[k:]
        weights = [1.0] + weights[k:]


def fedavg_flat(models: List[torch.Tensor], weights: List[float], out_dtype: torch.dtype = torch.float32) -> torch.Tensor:
    out = torch.empty(models[0].numel(), dtype=torch.float32, device=models[0].device)
    fedavg_into(models, weights, out, None)
    return out if out_dtype == torch.float32 else out.to(out_dtype)


# ------------------------------------------------------------------------ GEMM
def gemm_args(
    A: torch.Tensor, lda: int, B: torch.Tensor, ldb: int, M: int, N: int, K: int,
    bias: Optional[torch.Tensor] = None, bias_per_row: bool = True, relu: bool = False,
    mask: Optional[torch.Tensor] = None, ld_mask: int = 0,
    out_f32: Optional[torch.Tensor] = None, ld_f32: int = 0,
    out_bf16: Optional[torch.Tensor] = None, ld_bf16: int = 0,
    out_f32_t: Optional[torch.Tensor] = None, ld_f32_t: int = 0,
    out_bf16_t: Optional[torch.Tensor] = None, ld_bf16_t: int = 0,
    rowsum: Optional[torch.Tensor] = None, alpha: float = 1.0,
    split_k: int = 0, a_mn: bool = False, b_mn: bool = False,
) -> GemmArgs:  # fmt: skip
    """Describe D[M,N] = alpha * A[M,K] @ B[N,K]^T with its fused epilogue.

    ``a_mn`` / ``b_mn``: that operand is stored ``[K, ld]`` (M resp. N contiguous) instead.
    ``split_k`` (-1 = pick, n > 1 = n slices): K slices accumulate into a zero-filled ``out_f32``."""
    return GemmArgs(
        A.data_ptr(), lda, B.data_ptr(), ldb, M, N, K,
        0 if bias is None else bias.data_ptr(), int(bias_per_row), int(relu),
        0 if mask is None else mask.data_ptr(), ld_mask,
        0 if out_f32 is None else out_f32.data_ptr(), ld_f32,
        0 if out_bf16 is None else out_bf16.data_ptr(), ld_bf16,
        0 if out_f32_t is None else out_f32_t.data_ptr(), ld_f32_t,
        0 if out_bf16_t is None else out_bf16_t.data_ptr(), ld_bf16_t,
        0 if rowsum is None else rowsum.data_ptr(), alpha,
        split_k, int(a_mn) | (int(b_mn) << 1),
    )  # fmt: skip


def gemm_tn(A: torch.Tensor, *args, **kwargs) -> None:
    """One tcgen05 GEMM launch (see :func:`gemm_args` for the parameters)."""
    a = gemm_args(A, *args, **kwargs)
    _check(lib().p2pfl_gemm_tn(_dev(A), _stream(), C.byref(a)))
    _count()


def ce_args(logits, ld_logits: int, labels, B: int, Cn: int, dlogits, ld_d: int, dlogits_t, ld_dt: int, dbias, loss_sum) -> CeArgs:
    g = lambda t: 0 if t is None else t.data_ptr()  # noqa: E731
    return CeArgs(g(logits), ld_logits, g(labels), B, Cn, g(dlogits), ld_d, g(dlogits_t), ld_dt, g(dbias), g(loss_sum))


def gemm_chain(device: torch.device, gemms: Sequence[GemmArgs], links: Sequence[tuple], ce: Optional[CeArgs] = None,
               cluster: int = 1) -> None:
    """ONE launch for a list of dependent links: ``("gemm", i)`` runs ``gemms[i]``, ``("group", i, n)``
    runs the independent ``gemms[i:i+n]`` side by side, ``("ce", 0)`` the softmax-cross-entropy
    described by ``ce``; each link sees what the previous ones stored.  ``cluster`` CTAs (one
    thread-block cluster) share the tiles of every link."""
    arr = (GemmArgs * max(1, len(gemms)))(*gemms)
    kinds = (C.c_int32 * len(links))(*[1 if l[0] == "ce" else 0 for l in links])
    idxs = (C.c_int32 * len(links))(*[l[1] for l in links])
    cnts = (C.c_int32 * len(links))(*[l[2] if l[0] == "group" else 1 for l in links])
    dev = device.index if device.index is not None else torch.cuda.current_device()
    _check(
        lib().p2pfl_gemm_chain_cluster(
            dev, _stream(), arr, C.c_int(len(gemms)), kinds, idxs, cnts, C.c_int(len(links)),
            C.byref(ce) if ce is not None else None, C.c_int(cluster),
        )
    )
    _count()


_max_cluster: dict = {}


class MlpEpochArgs(C.Structure):
    """Mirror of ``p2pfl_mlp_epoch_args``."""

    _fields_ = [
        ("xs", C.c_void_p), ("ys", C.c_void_p), ("n_rows", C.c_int32), ("n_steps", C.c_int32),
        ("flat", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p), ("w16", C.c_void_p),
        ("grads_out", C.c_void_p), ("loss", C.c_void_p), ("step_ptr", C.c_void_p),
        ("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
    ]  # fmt: skip


MLP_EPOCH_DIMS = (784, 256, 128, 10)  # the architecture the epoch kernel is compiled for
MLP_EPOCH_BATCH = 32
_mlp_epoch_ok: dict = {}


def mlp_epoch_supported(device: torch.device) -> bool:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _mlp_epoch_ok:
        _mlp_epoch_ok[idx] = bool(lib().p2pfl_mlp_epoch_supported(C.c_int(idx)))
    return _mlp_epoch_ok[idx]


def mlp_epoch(xs, ys, n_rows: int, flat, m, v, w16, loss, step_ptr, lr: float, grads=None,
              beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8) -> None:  # fmt: skip
    """One training epoch of the 784-256-128-10 MLP over ``xs[:n_rows]`` (batches of 32 in row
    order) as a single cluster launch; updates ``flat`` / ``m`` / ``v`` / ``w16`` in place."""
    n_steps = (n_rows + MLP_EPOCH_BATCH - 1) // MLP_EPOCH_BATCH
    a = MlpEpochArgs(
        xs.data_ptr(), ys.data_ptr(), n_rows, n_steps, flat.data_ptr(), m.data_ptr(), v.data_ptr(), w16.data_ptr(),
        _p(grads).value, _p(loss).value, _p(step_ptr).value, lr, beta1, beta2, eps,
    )  # fmt: skip
    _check(lib().p2pfl_mlp_epoch(_dev(flat), _stream(), C.byref(a)))
    _count()


def gemm_chain_max_cluster(device: torch.device) -> int:
    """Largest cluster size (16, 8, 4, 2, 1) the chain kernel can be launched with on ``device``."""
    dev = device.index if device.index is not None else torch.cuda.current_device()
    if dev not in _max_cluster:
        _max_cluster[dev] = int(lib().p2pfl_gemm_chain_max_cluster(C.c_int(dev)))
    return _max_cluster[dev]


def gemm_tn_grouped(device: torch.device, problems: Sequence[GemmArgs]) -> None:
    """Up to four independent GEMMs in ONE launch."""
    arr = (GemmArgs * len(problems))(*problems)
    dev = device.index if device.index is not None else torch.cuda.current_device()
    _check(lib().p2pfl_gemm_tn_grouped(dev, _stream(), arr, C.c_int(len(problems))))
    _count()


# ---------------------------------------------------------------- data movement
def gather_rows(src, perm, n_rows: int, feat: int, batch: int, dst, dst_t=None, labels_src=None, labels_dst=None) -> None:
    _check(
        lib().p2pfl_gather_rows(
            _dev(src), _stream(), _p(src), _p(perm), C.c_int(n_rows), C.c_int(feat), C.c_int(batch),
            _p(dst), _p(dst_t), _p(labels_src), _p(labels_dst),
        )
    )  # fmt: skip
    _count()


# ------------------------------------------------------------ convolution path
def conv_out(size: int, k: int, stride: int, pad: int) -> int:
    return (size + 2 * pad - k) // stride + 1


def im2col(x, B: int, H: int, W: int, Cc: int, kh: int, kw: int, stride: int, pad: int, col, ld_col: int, col_t=None, ld_t: int = 0) -> None:
    _check(
        lib().p2pfl_im2col(
            _dev(x), _stream(), _p(x), C.c_int(B), C.c_int(H), C.c_int(W), C.c_int(Cc), C.c_int(kh), C.c_int(kw),
            C.c_int(stride), C.c_int(pad), _p(col), C.c_int(ld_col), _p(col_t), C.c_int64(ld_t),
        )
    )  # fmt: skip
    _count()


def col2im(dcol, ld_col: int, B: int, H: int, W: int, Cc: int, kh: int, kw: int, stride: int, pad: int, dx, extra=None) -> None:
    """dx = col2im(dcol) (+ extra, e.g. the gradient arriving through a residual branch)."""
    _check(
        lib().p2pfl_col2im_add(
            _dev(dcol), _stream(), _p(dcol), C.c_int(ld_col), C.c_int(B), C.c_int(H), C.c_int(W), C.c_int(Cc),
            C.c_int(kh), C.c_int(kw), C.c_int(stride), C.c_int(pad), _p(dx), _p(extra),
        )
    )  # fmt: skip
    _count()


def maxpool2_fwd(x, B: int, H: int, W: int, Cc: int, y, arg, flat=None, flat_t=None, ld_flat_t: int = 0) -> None:
    _check(
        lib().p2pfl_maxpool2_fwd(
            _dev(x), _stream(), _p(x), C.c_int(B), C.c_int(H), C.c_int(W), C.c_int(Cc), _p(y), _p(arg), _p(flat), _p(flat_t), C.c_int(ld_flat_t)
        )
    )
    _count()


def maxpool2_relu_bwd(dy_nhwc, dy_flat, arg, act, B: int, H: int, W: int, Cc: int, dx, dx_t=None, ld_t: int = 0, dbias=None) -> None:
    _check(
        lib().p2pfl_maxpool2_relu_bwd(
            _dev(act), _stream(), _p(dy_nhwc), _p(dy_flat), _p(arg), _p(act), C.c_int(B), C.c_int(H), C.c_int(W), C.c_int(Cc),
            _p(dx), _p(dx_t), C.c_int64(ld_t), _p(dbias),
        )
    )  # fmt: skip
    _count()


def pad_cast(src, rows: int, cols: int, dst, ld: int) -> None:
    _check(lib().p2pfl_pad_cast(_dev(src), _stream(), _p(src), C.c_int(rows), C.c_int(cols), _p(dst), C.c_int(ld)))
    _count()


def nchw_to_nhwc(x, B: int, Cc: int, H: int, W: int, y) -> None:
    _check(lib().p2pfl_nchw_to_nhwc(_dev(x), _stream(), _p(x), C.c_int(B), C.c_int(Cc), C.c_int(H), C.c_int(W), _p(y)))
    _count()


def im2col_kc(x, B: int, H: int, W: int, Cc: int, kh: int, kw: int, stride: int, pad: int, col, ld_col: int) -> None:
    """Tap-major im2col (K ordered (i, j, c)), 16-byte channel vectors; needs ``Cc % 8 == 0``."""
    _check(
        lib().p2pfl_im2col_kc(
            _dev(x), _stream(), _p(x), C.c_int(B), C.c_int(H), C.c_int(W), C.c_int(Cc), C.c_int(kh), C.c_int(kw),
            C.c_int(stride), C.c_int(pad), _p(col), C.c_int(ld_col),
        )
    )  # fmt: skip
    _count()


def col2im_kc(dcol, ld_col: int, B: int, H: int, W: int, Cc: int, kh: int, kw: int, stride: int, pad: int, dx, extra=None) -> None:
    _check(
        lib().p2pfl_col2im_kc(
            _dev(dcol), _stream(), _p(dcol), C.c_int(ld_col), C.c_int(B), C.c_int(H), C.c_int(W), C.c_int(Cc), C.c_int(kh),
            C.c_int(kw), C.c_int(stride), C.c_int(pad), _p(dx), _p(extra),
        )
    )  # fmt: skip
    _count()


def conv_pack(flat, packed, table, n_conv: int, max_elems: int) -> None:
    """fp32 PyTorch-order conv weights in ``flat`` -> bf16 packed ``[cout, ldk]`` copies, one launch."""
    _check(lib().p2pfl_conv_pack(_dev(flat), _stream(), _p(flat), _p(packed), _p(table), C.c_int(n_conv), C.c_int64(max_elems)))
    _count()


def conv_unpack(packed_f32, flat, table, n_conv: int, max_elems: int) -> None:
    """Packed-order fp32 weight gradients -> PyTorch order inside ``flat``, one launch."""
    _check(lib().p2pfl_conv_unpack(_dev(flat), _stream(), _p(packed_f32), _p(flat), _p(table), C.c_int(n_conv), C.c_int64(max_elems)))
    _count()


def bn_fwd(x, pixels: int, Cc: int, gamma, beta, running_mean, running_var, momentum: float, eps: float, relu: bool, residual, y, save_mean, save_rstd) -> None:
    _check(
        lib().p2pfl_bn_fwd(
            _dev(x), _stream(), _p(x), C.c_int64(pixels), C.c_int(Cc), _p(gamma), _p(beta), _p(running_mean), _p(running_var),
            C.c_float(momentum), C.c_float(eps), C.c_int(int(relu)), _p(residual), _p(y), _p(save_mean), _p(save_rstd),
        )
    )  # fmt: skip
    _count(3)


def bn_bwd(x, y, dy, pixels: int, Cc: int, gamma, save_mean, save_rstd, relu: bool, dx, d_residual, dgamma, dbeta, dx_t=None, ld_t: int = 0) -> None:
    _check(
        lib().p2pfl_bn_bwd_t(
            _dev(x), _stream(), _p(x), _p(y), _p(dy), C.c_int64(pixels), C.c_int(Cc), _p(gamma), _p(save_mean), _p(save_rstd),
            C.c_int(int(relu)), _p(dx), _p(dx_t), C.c_int64(ld_t), _p(d_residual), _p(dgamma), _p(dbeta),
        )
    )  # fmt: skip
    _count(2)


def bn_infer(x, pixels: int, Cc: int, gamma, beta, running_mean, running_var, eps: float, relu: bool, residual, y) -> None:
    _check(
        lib().p2pfl_bn_infer(
            _dev(x), _stream(), _p(x), C.c_int64(pixels), C.c_int(Cc), _p(gamma), _p(beta), _p(running_mean), _p(running_var),
            C.c_float(eps), C.c_int(int(relu)), _p(residual), _p(y),
        )
    )  # fmt: skip
    _count()


def avgpool_fwd(x, B: int, HW: int, Cc: int, y, y_t=None, ld_t: int = 0) -> None:
    _check(lib().p2pfl_avgpool_fwd(_dev(x), _stream(), _p(x), C.c_int(B), C.c_int(HW), C.c_int(Cc), _p(y), _p(y_t), C.c_int(ld_t)))
    _count()


def avgpool_bwd(dy, B: int, HW: int, Cc: int, dx) -> None:
    _check(lib().p2pfl_avgpool_bwd(_dev(dy), _stream(), _p(dy), C.c_int(B), C.c_int(HW), C.c_int(Cc), _p(dx)))
    _count()


# ---------------------------------------------------------------------- gossip
def make_push_table(inbox_ptrs: Sequence[int], flag_ptrs: Sequence[int], wire_bf16: bool) -> PushTable:
    if len(inbox_ptrs) > MAX_PEERS:
        raise ValueError(f"at most {MAX_PEERS} peers per push")
    t = PushTable()
    for i, (a, f) in enumerate(zip(inbox_ptrs, flag_ptrs)):
        t.inbox[i] = a
        t.flags[i] = f
    t.n_peers = len(inbox_ptrs)
    t.wire_bf16 = int(wire_bf16)
    return t


def make_reduce_table(
    slot_ptrs: Sequence[int], flag_ptrs: Sequence[int], weights: Sequence[float], self_pos: int, self_weight: float,
    wire_bf16: bool, round_self: bool = True,
) -> ReduceTable:  # fmt: skip
    if len(slot_ptrs) > MAX_PEERS:
        raise ValueError(f"at most {MAX_PEERS} contributors per reduce")
    t = ReduceTable()
    for i, (a, f, w) in enumerate(zip(slot_ptrs, flag_ptrs, weights)):
        t.slot[i] = a
        t.flags[i] = f
        t.weight[i] = w
    t.n_slots = len(slot_ptrs)
    t.self_pos = self_pos
    t.self_weight = self_weight
    t.wire_bf16 = int(wire_bf16)
    t.round_self = int(round_self)
    return t


def gossip_push(src: torch.Tensor, tile: int, seq: int, table: PushTable) -> None:
    _check(lib().p2pfl_gossip_push(_dev(src), _stream(), _p(src), C.c_uint64(src.numel()), C.c_uint32(tile), C.c_uint32(seq), C.byref(table)))
    _count()


def gossip_reduce(params, params_bf16, tile: int, seq: int, table: ReduceTable, abort_flag=None, timeout_cycles: int = 0, status=None) -> None:
    _check(
        lib().p2pfl_gossip_reduce(
            _dev(params), _stream(), _p(params), _p(params_bf16), C.c_uint64(params.numel()), C.c_uint32(tile), C.c_uint32(seq),
            C.byref(table), _p(abort_flag), C.c_uint64(timeout_cycles), _p(status),
        )
    )  # fmt: skip
    _count()


def gossip_fused(params, params_bf16, tile: int, seq: int, push: PushTable, reduce: ReduceTable, abort_flag=None,
                 timeout_cycles: int = 0, status=None, snapshot=None) -> None:  # fmt: skip
    """``snapshot`` (fp32, same length as ``params``): receives the pre-exchange model tile by tile."""
    _check(
        lib().p2pfl_gossip_fused_snap(
            _dev(params), _stream(), _p(params), _p(params_bf16), C.c_uint64(params.numel()), C.c_uint32(tile), C.c_uint32(seq),
            C.byref(push), C.byref(reduce), _p(abort_flag), C.c_uint64(timeout_cycles), _p(status), _p(snapshot),
        )
    )  # fmt: skip
    _count()


class GroupMember(C.Structure):
    """Mirror of ``p2pfl_group_member``."""

    _fields_ = [
        ("params", C.c_void_p),
        ("params_bf16", C.c_void_p),
        ("snapshot", C.c_void_p),
        ("abort_flag", C.c_void_p),
        ("status", C.c_void_p),
    ]


def make_group_member(params, params_bf16, snapshot, abort_flag, status) -> GroupMember:
    m = GroupMember()
    m.params, m.params_bf16, m.snapshot = params.data_ptr(), _p(params_bf16).value, _p(snapshot).value
    m.abort_flag, m.status = _p(abort_flag).value, status.data_ptr()
    return m


def gossip_fused_group(device: int, n: int, tile: int, seq: int, members_dev: torch.Tensor, push_dev: torch.Tensor,
                       reduce_dev: torch.Tensor, n_members: int, timeout_cycles: int = 0) -> None:  # fmt: skip
    """The exchange of every federated node of this process that trains on ``device``, as ONE
    cooperative launch on the current stream; the three tables are packed uint8 device tensors."""
    _check(
        lib().p2pfl_gossip_fused_group(
            C.c_int(device), _stream(), C.c_int(n_members), _p(members_dev), C.c_uint64(n), C.c_uint32(tile), C.c_uint32(seq),
            _p(push_dev), _p(reduce_dev), C.c_uint64(timeout_cycles),
        )
    )  # fmt: skip
    _count()


def gossip_fused_emulated(params: List[torch.Tensor], tile: int, seq: int, push_dev: torch.Tensor, reduce_dev: torch.Tensor, timeout_cycles: int, status: torch.Tensor) -> None:
    """Single-GPU emulation of ``len(params)`` ranks; ``push_dev`` / ``reduce_dev`` are
    uint8 device tensors holding the packed tables of all ranks."""
    r = len(params)
    ptrs = (C.c_void_p * r)(*[p.data_ptr() for p in params])
    _check(
        lib().p2pfl_gossip_fused_emulated(
            _dev(params[0]), _stream(), C.c_int(r), ptrs, C.c_uint64(params[0].numel()), C.c_uint32(tile), C.c_uint32(seq),
            _p(push_dev), _p(reduce_dev), C.c_uint64(timeout_cycles), _p(status),
        )
    )  # fmt: skip
    _count()


def gossip_push_bulk(src: torch.Tensor, seq: int, table: PushTable, begin: int = 0, end: Optional[int] = None) -> None:
    """Copy-engine push (cp.async.bulk) of ``src[begin:end]`` (whole 16384-element tiles) into every
    peer of ``table`` - inbox slots (fused exchange) or parameter regions (init / diffusion)."""
    n = src.numel()
    _check(
        lib().p2pfl_gossip_push_bulk(
            _dev(src), _stream(), _p(src), C.c_uint64(n), C.c_uint64(begin), C.c_uint64(n if end is None else end),
            C.c_uint32(seq), C.byref(table),
        )
    )  # fmt: skip
    _count()


def gossip_push_reduce_bulk(src: torch.Tensor, weight: float, table: PushTable) -> None:
    """Mode A: ``weight * src`` is ADDED into every peer's fp32 accumulator by the copy engine
    (cp.reduce.async.bulk); the table's ``flags`` are per-tile arrival counters."""
    _check(lib().p2pfl_gossip_push_reduce_bulk(_dev(src), _stream(), _p(src), C.c_uint64(src.numel()), C.c_float(weight), C.byref(table)))
    _count()


def gossip_finalize_acc(params, params_bf16, acc, counts, n_senders: int, self_weight: float, timeout_cycles: int = 0, status=None) -> None:
    """Receiver side of mode A: params = acc + self_weight * params once every sender's tile arrived."""
    _check(
        lib().p2pfl_gossip_finalize_acc(
            _dev(params), _stream(), _p(params), _p(params_bf16), _p(acc), _p(counts), C.c_uint64(params.numel()),
            C.c_int(n_senders), C.c_float(self_weight), C.c_uint64(timeout_cycles), _p(status),
        )
    )  # fmt: skip
    _count()


def tables_to_device(tables: Sequence[C.Structure], device: torch.device) -> torch.Tensor:
    raw = b"".join(bytes(t) for t in tables)
    return torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(device)


# --------------------------------------------------------------------- runtime
def rt_alloc(device: int, nbytes: int) -> int:
    out = C.c_void_p()
    _check(lib().p2pfl_rt_alloc(C.c_int(device), C.c_uint64(nbytes), C.byref(out)))
    return int(out.value)


def rt_free(device: int, ptr: int) -> None:
    _check(lib().p2pfl_rt_free(C.c_int(device), C.c_void_p(ptr)))


def rt_ipc_export(device: int, ptr: int) -> bytes:
    buf = (C.c_uint8 * 64)()
    _check(lib().p2pfl_rt_ipc_export(C.c_int(device), C.c_void_p(ptr), buf))
    return bytes(buf)


def rt_ipc_open(device: int, handle: bytes) -> int:
    out = C.c_void_p()
    buf = (C.c_uint8 * 64)(*handle)
    _check(lib().p2pfl_rt_ipc_open(C.c_int(device), buf, C.byref(out)))
    return int(out.value)


def rt_ipc_close(device: int, ptr: int) -> None:
    _check(lib().p2pfl_rt_ipc_close(C.c_int(device), C.c_void_p(ptr)))


_tls = threading.local()


_home_device: Optional[int] = None


def set_home_device(device: Optional[int]) -> None:
    """The GPU of this PROCESS (one process per GPU): threads of the framework make it their current
    device before their first CUDA call.  A new thread's current device is 0; its first runtime call
    then builds a context on GPU 0 - about 200 ms and half a gigabyte on every other rank, during which
    that thread (the gossiper that should deliver the greeting beat, the handler that should stash the
    initial model) is stuck.  Defaults to ``LOCAL_RANK`` when a launcher sets it."""
    global _home_device
    _home_device = device


def home_device() -> Optional[int]:
    if _home_device is not None:
        return _home_device
    try:
        return int(os.environ["LOCAL_RANK"])
    except (KeyError, ValueError):
        return None


def thread_uses_cuda() -> None:
    """Once per thread: current device = the process's home device (if it has one), and thread-local
    stream-capture mode (see ``p2pfl_rt_thread_capture_mode_local``).
    No-op without the kernel library or without a CUDA device."""
    if getattr(_tls, "capture_mode_set", False):
        return
    _tls.capture_mode_set = True
    try:
        if available() and torch.cuda.is_available():
            dev = home_device()
            if dev is not None and 0 <= dev < torch.cuda.device_count():
                torch.cuda.set_device(dev)
            lib().p2pfl_rt_thread_capture_mode_local()
    except Exception:
        pass


_clock_hz: dict = {}


def rt_sm_clock_hz(device: int) -> float:
    """Peak SM clock of ``device`` in Hz (cached)."""
    hz = _clock_hz.get(device)
    if hz is None:
        khz = C.c_int(0)
        _check(lib().p2pfl_rt_sm_clock_khz(C.c_int(device), C.byref(khz)))
        hz = _clock_hz[device] = float(khz.value) * 1e3 if khz.value > 0 else 1.9e9
    return hz


WATCH_DONE, WATCH_CANCELLED, WATCH_TIMEOUT, WATCH_SLICE = 0, 1, 2, 3


def rt_watch(device: int, done_event: Any, abort_word: Any = None, cancel: Any = None, timeout_s: float = -1.0,
             slice_s: float = 0.0, grace_s: float = 30.0) -> Tuple[int, float]:  # fmt: skip
    """Native wait for ``done_event`` (a recorded ``torch.cuda.Event``): see ``p2pfl_rt_watch`` in the
    header.  ``abort_word``: device int32 tensor; ``cancel``: a ``ctypes.c_int32`` other threads may
    set.  The interpreter lock is released for the whole call.  Returns (verdict, seconds waited)."""
    waited = C.c_double(0.0)
    rc = lib().p2pfl_rt_watch(
        C.c_int(device), C.c_void_p(int(done_event.cuda_event)), _p(abort_word),
        C.byref(cancel) if cancel is not None else None, C.c_double(timeout_s), C.c_double(slice_s), C.c_double(grace_s),
        C.byref(waited),
    )  # fmt: skip
    if rc not in (WATCH_DONE, WATCH_CANCELLED, WATCH_TIMEOUT, WATCH_SLICE):
        _check(rc if rc != 0 else -1)
    return rc, float(waited.value)


def rt_stream_create(device: int) -> int:
    out = C.c_void_p()
    _check(lib().p2pfl_rt_stream_create(C.c_int(device), C.byref(out)))
    return int(out.value)


def rt_stream_destroy(device: int, stream: int) -> None:
    lib().p2pfl_rt_stream_destroy(C.c_int(device), C.c_void_p(stream))


def rt_enable_peer(device: int, peer: int) -> None:
    _check(lib().p2pfl_rt_enable_peer(C.c_int(device), C.c_int(peer)))


def rt_can_access_peer(device: int, peer: int) -> bool:
    out = C.c_int()
    _check(lib().p2pfl_rt_can_access_peer(C.c_int(device), C.c_int(peer), C.byref(out)))
    return bool(out.value)
