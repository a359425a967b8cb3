"""Fused sm_100a train step of the CIFAR-shaped ResNet-18 (BASELINE.json config 3).

Every convolution is a GEMM on the tcgen05 kernel over a tap-major column matrix
(``im2col_kc``: K ordered (i, j, c), so NHWC activations move as 16-byte channel vectors).
The backward GEMMs read the SAME buffers as MN-major tensor-core operands, so nothing is
transposed: weight gradient ``dZ[pix, Cout]^T . col[pix, K]`` (split over K = pixel count, the
slices accumulate atomically), data gradient ``dZ . W[Cout, K]`` + ``col2im_kc``.  Every
BatchNorm (+ residual add, + ReLU) is the coalesced stats / apply kernel pair, the head is
global average pool + a swap-AB Linear + the fused softmax-cross-entropy.  Activations are
NHWC bf16; parameters, BatchNorm running statistics, gradients and Adam state live in the
learner's flat arena (PyTorch weight order - the tap-major bf16 copies are repacked by one
launch per step), so ``state_dict`` and the gossip data plane are unchanged.

Not part of the reference (it ships only the MLP and the small CNN); built for the
"CIFAR-10-shaped ResNet-18, ring, bf16" configuration.
"""

from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch
from torch import nn

from p2pfl_b200 import ops


def _pad8(x: int) -> int:
    return (x + 7) // 8 * 8


class _Conv:
    """One convolution + the BatchNorm that follows it."""

    def __init__(self, plan: "ResNetPlan", name: str, bn: str, conv: nn.Conv2d, H: int) -> None:
        lay = plan.lr.layout.by_name
        self.name, self.bn = name, bn
        self.cin, self.cout = conv.in_channels, conv.out_channels
        self.k, self.stride, self.pad = conv.kernel_size[0], conv.stride[0], conv.padding[0]
        self.Hin, self.Hout = H, ops.conv_out(H, self.k, self.stride, self.pad)
        self.K = self.cin * self.k * self.k
        self.ldk = _pad8(self.K)
        self.w_off = lay[name + ".weight"].offset
        self.g_off, self.b_off = lay[bn + ".weight"].offset, lay[bn + ".bias"].offset
        self.rm_off, self.rv_off = lay[bn + ".running_mean"].offset, lay[bn + ".running_var"].offset
        self.kc = self.cin % 8 == 0  # tap-major K order; the 3-channel stem keeps PyTorch's (c, i, j)
        self.pk_off = -1  # offset of this conv in the packed weight / packed gradient buffers


class _Block:
    def __init__(self, c1: _Conv, c2: _Conv, ds: Optional[_Conv]) -> None:
        self.c1, self.c2, self.ds = c1, c2, ds


class ResNetPlan:
    EVAL_CHUNK = 250
    BN_MOMENTUM, BN_EPS = 0.1, 1e-5
    nhwc_input = True  # the learner permutes the dataset to NHWC once, at upload

    def __init__(self, learner: Any) -> None:
        self.lr = learner
        m = learner.model
        self.ok = False
        need = ("conv1", "bn1", "layer1", "layer2", "layer3", "layer4", "fc")
        if not all(hasattr(m, n) for n in need) or not isinstance(m.fc, nn.Linear):
            return
        shape = getattr(m, "in_shape", (3, 32, 32))
        self.cin, self.H = int(shape[0]), int(shape[1])
        self.feat = self.cin * self.H * self.H
        H = self.H
        self.stem = _Conv(self, "conv1", "bn1", m.conv1, H)
        H = self.stem.Hout
        self.blocks: List[_Block] = []
        for ln in ("layer1", "layer2", "layer3", "layer4"):
            for bi, blk in enumerate(getattr(m, ln)):
                p = f"{ln}.{bi}"
                c1 = _Conv(self, p + ".conv1", p + ".bn1", blk.conv1, H)
                c2 = _Conv(self, p + ".conv2", p + ".bn2", blk.conv2, c1.Hout)
                ds = None
                if blk.downsample is not None:
                    ds = _Conv(self, p + ".downsample.0", p + ".downsample.1", blk.downsample[0], H)
                self.blocks.append(_Block(c1, c2, ds))
                H = c2.Hout
        self.convs: List[_Conv] = [self.stem]
        for b in self.blocks:
            self.convs += [b.c1, b.c2] + ([b.ds] if b.ds is not None else [])
        if any(c.cout % 8 for c in self.convs):
            return
        self.Hlast, self.clast = H, self.blocks[-1].c2.cout
        self.ncls = m.fc.out_features
        if m.fc.in_features != self.clast or m.fc.bias is None:
            return
        lay = learner.layout.by_name
        self.fcw_off, self.fcb_off = lay["fc.weight"].offset, lay["fc.bias"].offset
        dev = learner.device
        # packed bf16 conv weights [cout, ldk] (and, same offsets, the fp32 gradients the wgrad
        # GEMMs accumulate); `table` drives the one-launch pack / unpack kernels
        off, rows = 0, []
        for c in self.convs:
            c.pk_off = off
            rows.append([c.w_off, off, c.cout, c.cin, c.k * c.k, c.ldk, int(c.kc), 0])
            off += (c.cout * c.ldk + 63) // 64 * 64
        self.wk = torch.zeros(off, dtype=torch.bfloat16, device=dev)
        self.gw = torch.zeros(off, dtype=torch.float32, device=dev)
        self.table = torch.tensor(rows, dtype=torch.int64, device=dev)
        self.max_w = max(c.cout * c.K for c in self.convs)
        # the head's W^T [clast, ncls] (dgrad operand of the swap-AB Linear)
        self.t_fc = 0
        self.tdescs = [(self.fcw_off, 0, self.ncls, self.clast, _pad8(self.ncls))]
        self.wT16 = torch.zeros(self.clast * _pad8(self.ncls), dtype=torch.bfloat16, device=dev)
        self.ok = True

    # ------------------------------------------------------------------ copies
    def refresh(self) -> None:
        L = self.lr
        ops.conv_pack(L.flat, self.wk, self.table, len(self.convs), self.max_w)
        ops.transpose_cast(L.flat, self.wT16, self.tdescs)

    def _w(self, c: _Conv) -> torch.Tensor:
        return self.wk[c.pk_off :]

    def _im2col(self, c: _Conv, x: torch.Tensor, bsz: int, col: torch.Tensor) -> None:
        if c.kc:
            ops.im2col_kc(x, bsz, c.Hin, c.Hin, c.cin, c.k, c.k, c.stride, c.pad, col, c.ldk)
        else:
            ops.im2col(x, bsz, c.Hin, c.Hin, c.cin, c.k, c.k, c.stride, c.pad, col, c.ldk)

    # ----------------------------------------------------------------- buffers
    def alloc(self, B: int, train: bool) -> Dict[str, Any]:
        dev = self.lr.device
        z = lambda *s, dt=torch.bfloat16: torch.zeros(*s, dtype=dt, device=dev)  # noqa: E731
        a: Dict[str, Any] = {"B": B, "ldB": _pad8(B), "ld_logits": _pad8(self.ncls), "conv": {}}
        max_act = max_col = 0
        for c in self.convs:
            pix = B * c.Hout * c.Hout
            d = {"col": z(pix, c.ldk), "z": z(pix, c.cout), "y": z(pix, c.cout)}
            if train:
                d["sm"] = z(c.cout, dt=torch.float32)
                d["sr"] = z(c.cout, dt=torch.float32)
            a["conv"][c.name] = d
            max_act = max(max_act, pix * c.cout, B * c.Hin * c.Hin * c.cin)
            max_col = max(max_col, pix * c.ldk)
        a["f"] = z(B, self.clast)
        a["logits"] = z(B, _pad8(self.ncls), dt=torch.float32)
        if train:
            a["fT"] = z(self.clast, a["ldB"])
            a["dZ"] = z(B, _pad8(self.ncls))
            a["dZT"] = z(self.ncls, a["ldB"])
            a["df"] = z(B, self.clast)
            # gradient scratch, sized for the widest layer and reused down the network
            for k in ("g0", "g1", "g2", "g3", "dz"):
                a[k] = z(max_act)
            a["dcol"] = z(max_col)
        return a

    # ----------------------------------------------------------------- forward
    def _conv_bn(self, c: _Conv, x: torch.Tensor, bsz: int, a: Dict[str, Any], train: bool, relu: bool,
                 residual: Optional[torch.Tensor]) -> torch.Tensor:
        """y = [relu](bn(conv(x)) [+ residual]); x NHWC bf16 [bsz, Hin, Hin, cin]."""
        L, d = self.lr, a["conv"][c.name]
        pix = bsz * c.Hout * c.Hout
        self._im2col(c, x, bsz, d["col"])
        ops.gemm_tn(d["col"], c.ldk, self._w(c), c.ldk, pix, c.cout, c.K, out_bf16=d["z"], ld_bf16=c.cout)
        f = L.flat
        if train:
            ops.bn_fwd(d["z"], pix, c.cout, f[c.g_off :], f[c.b_off :], f[c.rm_off :], f[c.rv_off :], self.BN_MOMENTUM,
                       self.BN_EPS, relu, residual, d["y"], d["sm"], d["sr"])  # fmt: skip
        else:
            ops.bn_infer(d["z"], pix, c.cout, f[c.g_off :], f[c.b_off :], f[c.rm_off :], f[c.rv_off :], self.BN_EPS,
                         relu, residual, d["y"])  # fmt: skip
        return d["y"]

    def forward(self, x16: torch.Tensor, bsz: int, a: Dict[str, Any], train: bool) -> None:
        """x16: bf16 NHWC input rows [bsz, H*H*cin]."""
        L = self.lr
        cur = self._conv_bn(self.stem, x16, bsz, a, train, True, None)
        for b in self.blocks:
            y1 = self._conv_bn(b.c1, cur, bsz, a, train, True, None)
            res = cur if b.ds is None else self._conv_bn(b.ds, cur, bsz, a, train, False, None)
            cur = self._conv_bn(b.c2, y1, bsz, a, train, True, res)
        hw = self.Hlast * self.Hlast
        ops.avgpool_fwd(cur, bsz, hw, self.clast, a["f"], a.get("fT") if train else None, a["ldB"])
        ops.gemm_tn(L.w16[self.fcw_off :], self.clast, a["f"], self.clast, self.ncls, bsz, self.clast,
                    bias=L.flat[self.fcb_off :], bias_per_row=True, out_f32_t=a["logits"], ld_f32_t=a["ld_logits"])  # fmt: skip

    def forward_eval(self, x16: torch.Tensor, n: int, a: Dict[str, Any], logits_out: torch.Tensor) -> None:
        chunk = a["B"]
        for c0 in range(0, n, chunk):
            bsz = min(chunk, n - c0)
            self.forward(x16[c0:], bsz, a, train=False)
            logits_out[c0 : c0 + bsz].copy_(a["logits"][:bsz])

    # ---------------------------------------------------------------- backward
    def _bn_conv_bwd(self, c: _Conv, x_in: torch.Tensor, dy: torch.Tensor, bsz: int, a: Dict[str, Any], relu: bool,
                     d_res: Optional[torch.Tensor], dx_out: Optional[torch.Tensor], extra: Optional[torch.Tensor]) -> None:
        """Backward through [relu] <- bn <- conv.  ``dy``: gradient w.r.t. the block's output y;
        ``d_res`` (optional) receives the ReLU-masked gradient that flows into the residual branch;
        ``dx_out`` (optional) receives the gradient w.r.t. the conv input (+ ``extra``)."""
        L, d, g = self.lr, a["conv"][c.name], self.lr.grads
        pix = bsz * c.Hout * c.Hout
        f = L.flat
        ops.bn_bwd(d["z"], d["y"], dy, pix, c.cout, f[c.g_off :], d["sm"], d["sr"], relu, a["dz"], d_res,
                   g[c.g_off :], g[c.b_off :])  # fmt: skip
        # dW[cout, K] = dz[pix, cout]^T . col[pix, K]: both operands MN-major, the pixel axis split
        probs = [ops.gemm_args(a["dz"], c.cout, d["col"], c.ldk, c.cout, c.K, pix, out_f32=self.gw[c.pk_off :], ld_f32=c.ldk,
                               split_k=-1, a_mn=True, b_mn=True)]  # fmt: skip
        if dx_out is not None:
            # dcol[pix, K] = dz[pix, cout] . W[cout, K]: the packed weight is the MN-major B operand
            probs.append(ops.gemm_args(a["dz"], c.cout, self._w(c), c.ldk, pix, c.K, c.cout, out_bf16=a["dcol"],
                                       ld_bf16=c.ldk, b_mn=True))  # fmt: skip
        if len(probs) == 2 and ((pix + 127) // 128) * ((c.K + 255) // 256) >= 296:
            # a data gradient with many more tiles than SMs runs on the persistent kernel (its
            # epilogue overlaps the next tile); the split-K weight gradient gets its own launch
            for pr in probs:
                ops.gemm_tn_grouped(L.device, [pr])
        else:
            ops.gemm_tn_grouped(L.device, probs)
        if dx_out is not None:
            ops.col2im_kc(a["dcol"], c.ldk, bsz, c.Hin, c.Hin, c.cin, c.k, c.k, c.stride, c.pad, dx_out, extra)

    def train_step(self, x16, y, bsz: int, a: Dict[str, Any], loss_slot, step_offset: int, lr: float) -> None:
        L, g = self.lr, self.lr.grads
        ldB = a["ldB"]
        self.gw.zero_()  # the split-K weight gradients accumulate
        self.forward(x16, bsz, a, train=True)
        ops.softmax_xent(a["logits"], a["ld_logits"], y, bsz, self.ncls, a["dZ"], _pad8(self.ncls), a["dZT"], ldB,
                         g[self.fcb_off :], loss_slot)  # fmt: skip
        # head: wgrad + dgrad (gradient w.r.t. the pooled features)
        ops.gemm_tn_grouped(L.device, [
            ops.gemm_args(a["dZT"], ldB, a["fT"], ldB, self.ncls, self.clast, bsz, out_f32=g[self.fcw_off :], ld_f32=self.clast),
            ops.gemm_args(self.wT16[self.t_fc :], _pad8(self.ncls), a["dZ"], _pad8(self.ncls), self.clast, bsz, self.ncls,
                          out_bf16_t=a["df"], ld_bf16_t=self.clast),
        ])  # fmt: skip
        hw = self.Hlast * self.Hlast
        dcur, spare = a["g0"], a["g1"]
        ops.avgpool_bwd(a["df"], bsz, hw, self.clast, dcur)
        for bi in range(len(self.blocks) - 1, -1, -1):
            b = self.blocks[bi]
            x_in = a["conv"][self.blocks[bi - 1].c2.name]["y"] if bi > 0 else a["conv"][self.stem.name]["y"]
            y1 = a["conv"][b.c1.name]["y"]
            # out = relu(bn2(conv2(y1)) + res): d_res = masked dcur, dy1 = gradient into y1
            self._bn_conv_bwd(b.c2, y1, dcur, bsz, a, True, a["g2"], a["g3"], None)
            if b.ds is None:
                # dx = (conv1 path) + d_res
                self._bn_conv_bwd(b.c1, x_in, a["g3"], bsz, a, True, None, spare, a["g2"])
            else:
                # residual branch: bn_d <- conv_d; its input gradient is added to the conv1 path's
                self._bn_conv_bwd(b.ds, x_in, a["g2"], bsz, a, False, None, dcur, None)
                self._bn_conv_bwd(b.c1, x_in, a["g3"], bsz, a, True, None, spare, dcur)
            dcur, spare = spare, dcur
        self._bn_conv_bwd(self.stem, x16, dcur, bsz, a, True, None, None, None)
        ops.conv_unpack(self.gw, g, self.table, len(self.convs), self.max_w)
        ops.adam_step(L.flat, g, L.adam_m, L.adam_v, L.w16, lr, L.adam_step, step_offset)
        self.refresh()
