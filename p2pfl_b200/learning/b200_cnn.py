"""Fused sm_100a train step of the MNIST-shaped CNN (reference: mnist_examples/models/cnn.py:55-91).

    conv5x5(1->32,'same') ReLU pool2  conv5x5(32->64,'same') ReLU pool2  fc(3136->2048) ReLU  fc(2048->10)

Every multiply-accumulate runs on the tcgen05 GEMM (``ops.gemm_tn``):

* convolutions are implicit GEMMs: ``im2col`` builds ``col[pixels, Cin*25]`` (and its
  transpose) from NHWC bf16 activations, forward is ``col . W^T`` with bias + ReLU fused
  in the epilogue, the weight gradient is ``dOut^T . col`` (K = pixel count), the data
  gradient is ``dOut . W`` followed by the gather-form ``col2im``;
* ReLU + 2x2 max-pool backward, the conv bias gradients and the layout changes are one
  pass (``maxpool2_relu_bwd``); pool 2 writes its output directly in the NCHW-flatten
  order ``nn.Linear`` expects, so ``state_dict`` stays interchangeable with the reference;
* the two Linear layers use the same swap-AB scheme as the MLP path;
* each layer's wgrad + dgrad pair is ONE grouped launch; Adam also refreshes the bf16
  transposed weight copies.

The object owns the step's buffers; ``B200Learner`` captures the epoch as a CUDA graph.
"""

from __future__ import annotations

from typing import Any, Dict

import torch
from torch import nn

from p2pfl_b200 import ops


def _pad8(x: int) -> int:
    return (x + 7) // 8 * 8


class CnnPlan:
    EVAL_CHUNK = 250

    def __init__(self, learner: Any) -> None:
        self.lr = learner
        m = learner.model
        self.ok = False
        if not all(hasattr(m, n) for n in ("conv1", "conv2", "l1", "l2")):
            return
        c1, c2, l1, l2 = m.conv1, m.conv2, m.l1, m.l2
        if not (isinstance(c1, nn.Conv2d) and isinstance(c2, nn.Conv2d) and isinstance(l1, nn.Linear) and isinstance(l2, nn.Linear)):
            return
        if c1.kernel_size != c2.kernel_size or c1.stride != (1, 1) or c2.stride != (1, 1):
            return
        self.k = c1.kernel_size[0]
        self.pad = self.k // 2
        self.cin, self.c1, self.c2 = c1.in_channels, c1.out_channels, c2.out_channels
        self.H = int(getattr(m, "in_shape", (1, 28, 28))[1])
        self.feat = self.cin * self.H * self.H
        self.H2, self.H4 = self.H // 2, self.H // 4
        self.f_in = self.c2 * self.H4 * self.H4
        if l1.in_features != self.f_in or l2.in_features != l1.out_features or (self.c1 * self.k * self.k) % 8:
            return
        self.hid, self.ncls = l1.out_features, l2.out_features
        lay = learner.layout.by_name
        self.off = {n: lay[n].offset for n in (
            "conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "l1.weight", "l1.bias", "l2.weight", "l2.bias")}  # fmt: skip
        self.K1, self.K2 = self.cin * self.k * self.k, self.c1 * self.k * self.k
        dev = learner.device
        # bf16 weight copies that are not plain views of w16
        self.w1p = torch.zeros(self.c1, _pad8(self.K1), dtype=torch.bfloat16, device=dev)  # padded conv1 weight
        o2 = 0
        self.t_w2 = o2  # W2^T  [K2, c2]
        o2 += self.K2 * _pad8(self.c2)
        self.t_l1 = o2  # Wl1^T [f_in, hid]
        o2 += self.f_in * _pad8(self.hid)
        self.t_l2 = o2  # Wl2^T [hid, ncls]
        o2 += self.hid * _pad8(self.ncls)
        self.wT16 = torch.zeros(o2, dtype=torch.bfloat16, device=dev)
        self.tdescs = [
            (self.off["conv2.weight"], self.t_w2, self.c2, self.K2, _pad8(self.c2)),
            (self.off["l1.weight"], self.t_l1, self.hid, self.f_in, _pad8(self.hid)),
            (self.off["l2.weight"], self.t_l2, self.ncls, self.hid, _pad8(self.ncls)),
        ]
        self.ok = True

    # ------------------------------------------------------------------ copies
    def refresh(self) -> None:
        """fp32 master changed: rebuild the bf16 copies that are not views of ``w16``."""
        L = self.lr
        ops.transpose_cast(L.flat, self.wT16, self.tdescs)
        ops.pad_cast(L.flat[self.off["conv1.weight"] :], self.c1, self.K1, self.w1p, _pad8(self.K1))

    # ----------------------------------------------------------------- buffers
    def alloc(self, B: int, train: bool) -> Dict[str, Any]:
        dev = self.lr.device
        z = lambda *s, dt=torch.bfloat16: torch.zeros(*s, dtype=dt, device=dev)  # noqa: E731
        H, H2, H4 = self.H, self.H2, self.H4
        p1, p2 = B * H * H, B * H2 * H2
        ldB = _pad8(B)
        a: Dict[str, Any] = {"B": B, "ldB": ldB, "ld_logits": _pad8(self.ncls)}
        a["col1"] = z(p1, _pad8(self.K1))
        a["a1"] = z(p1, self.c1)
        a["arg1"] = z(B * H2 * H2 * self.c1, dt=torch.uint8)
        a["p1"] = z(p2, self.c1)
        a["col2"] = z(p2, self.K2)
        a["a2"] = z(p2, self.c2)
        a["arg2"] = z(B * H4 * H4 * self.c2, dt=torch.uint8)
        a["f"] = z(B, self.f_in)
        a["H1"] = z(B, _pad8(self.hid))
        a["logits"] = z(B, _pad8(self.ncls), dt=torch.float32)
        if train:
            a["col1T"] = z(self.K1, p1)
            a["col2T"] = z(self.K2, p2)
            a["fT"] = z(self.f_in, ldB)
            a["HT1"] = z(self.hid, ldB)
            a["dZ2"] = z(B, _pad8(self.ncls))
            a["dZ2T"] = z(self.ncls, ldB)
            a["dZ1"] = z(B, _pad8(self.hid))
            a["dZ1T"] = z(self.hid, ldB)
            a["df"] = z(B, self.f_in)
            a["dO2"] = z(p2, self.c2)
            a["dO2T"] = z(self.c2, p2)
            a["dcol2"] = z(p2, self.K2)
            a["dP1"] = z(p2, self.c1)
            a["dO1"] = z(p1, self.c1)
            a["dO1T"] = z(self.c1, p1)
        return a

    # ----------------------------------------------------------------- forward
    def forward(self, x16: torch.Tensor, bsz: int, a: Dict[str, Any], train: bool) -> None:
        """x16: bf16 [bsz, cin*H*H]; for cin == 1 NCHW and NHWC coincide."""
        L, k, pad, H, H2 = self.lr, self.k, self.pad, self.H, self.H2
        p1, p2 = bsz * H * H, bsz * H2 * H2
        ld1 = _pad8(self.K1)
        p1_alloc, p2_alloc = a["B"] * H * H, a["B"] * H2 * H2
        ops.im2col(x16, bsz, H, H, self.cin, k, k, 1, pad, a["col1"], ld1, a.get("col1T") if train else None, p1_alloc)
        ops.gemm_tn(a["col1"], ld1, self.w1p, ld1, p1, self.c1, self.K1, bias=L.flat[self.off["conv1.bias"] :],
                    bias_per_row=False, relu=True, out_bf16=a["a1"], ld_bf16=self.c1)  # fmt: skip
        ops.maxpool2_fwd(a["a1"], bsz, H, H, self.c1, a["p1"], a["arg1"])
        ops.im2col(a["p1"], bsz, H2, H2, self.c1, k, k, 1, pad, a["col2"], self.K2, a.get("col2T") if train else None, p2_alloc)
        ops.gemm_tn(a["col2"], self.K2, L.w16[self.off["conv2.weight"] :], self.K2, p2, self.c2, self.K2,
                    bias=L.flat[self.off["conv2.bias"] :], bias_per_row=False, relu=True, out_bf16=a["a2"], ld_bf16=self.c2)  # fmt: skip
        ops.maxpool2_fwd(a["a2"], bsz, H2, H2, self.c2, None, a["arg2"], a["f"], a.get("fT") if train else None, a["ldB"])
        # fc1 / fc2 in swap-AB form (out-features on the 128-row MMA dimension)
        ops.gemm_tn(L.w16[self.off["l1.weight"] :], self.f_in, a["f"], self.f_in, self.hid, bsz, self.f_in,
                    bias=L.flat[self.off["l1.bias"] :], bias_per_row=True, relu=True,
                    out_bf16=a.get("HT1") if train else None, ld_bf16=a["ldB"], out_bf16_t=a["H1"], ld_bf16_t=_pad8(self.hid))  # fmt: skip
        ops.gemm_tn(L.w16[self.off["l2.weight"] :], self.hid, a["H1"], _pad8(self.hid), self.ncls, bsz, self.hid,
                    bias=L.flat[self.off["l2.bias"] :], bias_per_row=True, out_f32_t=a["logits"], ld_f32_t=a["ld_logits"])  # fmt: skip

    def forward_eval(self, x16: torch.Tensor, n: int, a: Dict[str, Any], logits_out: torch.Tensor) -> None:
        """Evaluation forward in chunks (the im2col operands of a whole split would be hundreds of MB)."""
        chunk = a["B"]
        for c0 in range(0, n, chunk):
            bsz = min(chunk, n - c0)
            self.forward(x16[c0:], bsz, a, train=False)
            logits_out[c0 : c0 + bsz].copy_(a["logits"][:bsz])

    # -------------------------------------------------------------------- step
    def train_step(self, x16, y, bsz: int, a: Dict[str, Any], loss_slot, step_offset: int, lr: float) -> None:
        L, k, pad, H, H2 = self.lr, self.k, self.pad, self.H, self.H2
        p1, p2 = bsz * H * H, bsz * H2 * H2
        p1_alloc, p2_alloc = a["B"] * H * H, a["B"] * H2 * H2
        ldB, g = a["ldB"], L.grads
        g.zero_()  # conv bias gradients are accumulated with atomics
        self.forward(x16, bsz, a, train=True)
        ops.softmax_xent(a["logits"], a["ld_logits"], y, bsz, self.ncls, a["dZ2"], _pad8(self.ncls), a["dZ2T"], ldB,
                         g[self.off["l2.bias"] :], loss_slot)  # fmt: skip
        # fc2: wgrad + dgrad (ReLU mask of fc1's output, row sums = fc1 bias gradient)
        ops.gemm_tn_grouped(L.device, [
            ops.gemm_args(a["dZ2T"], ldB, a["HT1"], ldB, self.ncls, self.hid, bsz, out_f32=g[self.off["l2.weight"] :], ld_f32=self.hid),
            ops.gemm_args(self.wT16[self.t_l2 :], _pad8(self.ncls), a["dZ2"], _pad8(self.ncls), self.hid, bsz, self.ncls,
                          mask=a["HT1"], ld_mask=ldB, out_bf16=a["dZ1T"], ld_bf16=ldB, out_bf16_t=a["dZ1"],
                          ld_bf16_t=_pad8(self.hid), rowsum=g[self.off["l1.bias"] :]),
        ])  # fmt: skip
        # fc1: wgrad + dgrad (gradient w.r.t. the flattened pool-2 output, NCHW-flatten order)
        ops.gemm_tn_grouped(L.device, [
            ops.gemm_args(a["dZ1T"], ldB, a["fT"], ldB, self.hid, self.f_in, bsz, out_f32=g[self.off["l1.weight"] :], ld_f32=self.f_in),
            ops.gemm_args(self.wT16[self.t_l1 :], _pad8(self.hid), a["dZ1"], _pad8(self.hid), self.f_in, bsz, self.hid,
                          out_bf16_t=a["df"], ld_bf16_t=self.f_in),
        ])  # fmt: skip
        ops.maxpool2_relu_bwd(None, a["df"], a["arg2"], a["a2"], bsz, H2, H2, self.c2, a["dO2"], a["dO2T"], p2_alloc,
                              g[self.off["conv2.bias"] :])  # fmt: skip
        # conv2: wgrad (K = pixels) + dgrad
        ops.gemm_tn_grouped(L.device, [
            ops.gemm_args(a["dO2T"], p2_alloc, a["col2T"], p2_alloc, self.c2, self.K2, p2, out_f32=g[self.off["conv2.weight"] :], ld_f32=self.K2),
            ops.gemm_args(a["dO2"], self.c2, self.wT16[self.t_w2 :], _pad8(self.c2), p2, self.K2, self.c2, out_bf16=a["dcol2"], ld_bf16=self.K2),
        ])  # fmt: skip
        ops.col2im(a["dcol2"], self.K2, bsz, H2, H2, self.c1, k, k, 1, pad, a["dP1"])
        ops.maxpool2_relu_bwd(a["dP1"], None, a["arg1"], a["a1"], bsz, H, H, self.c1, a["dO1"], a["dO1T"], p1_alloc,
                              g[self.off["conv1.bias"] :])  # fmt: skip
        ops.gemm_tn(a["dO1T"], p1_alloc, a["col1T"], p1_alloc, self.c1, self.K1, p1, out_f32=g[self.off["conv1.weight"] :], ld_f32=self.K1)
        ops.adam_step_t(L.flat, g, L.adam_m, L.adam_v, L.w16, lr, L.adam_step, step_offset, self.wT16, self.tdescs)
        ops.pad_cast(L.flat[self.off["conv1.weight"] :], self.c1, self.K1, self.w1p, _pad8(self.K1))
