"""Fused sm_100a learner: the local train step as hand-written kernels.

Replaces what the reference does inside ``Trainer.fit`` / ``Trainer.test``
(p2pfl/learning/pytorch/lightning_learner.py:180-236) for the
MNIST-shaped MLP family (mlp.py:53-69):

* parameters live in a peer-mappable device arena (fp32 master) with a bf16
  working copy and a bf16 transposed copy of each weight for the dgrad GEMM;
* the 784-256-128-10 MLP at batch 32 (the benchmark's model) trains a WHOLE
  EPOCH in one launch of the cluster kernel ``ops.mlp_epoch``
  (csrc/mlp_cluster.cu, ``Settings.MLP_EPOCH_KERNEL``): 16 CTAs keep the bf16
  weights in shared memory and the fp32 master weights and Adam moments of
  layer 1 in tensor memory across steps, hand activations to each other
  through distributed shared memory, and apply Adam where the weight slice
  lives - 11.7 us per step;
* other MLP shapes / batch sizes run every Linear forward / dgrad / wgrad as the
  tcgen05 GEMM with bias + ReLU + ReLU-mask + bias-gradient fused in its
  epilogue: ONE thread-block cluster walks all of them plus the loss as dependent
  links of one launch (``ops.gemm_chain(..., cluster=16)``,
  ``Settings.MLP_CHAIN_CLUSTER``) and Adam over the whole arena (+ bf16 refresh
  + transposed copies) is the second launch of the step; ``MLP_CHAIN_CLUSTER = 0``
  runs one launch per link instead (swap-AB GEMMs, a fused log-softmax +
  cross-entropy kernel);
* the CNN and ResNet-18 families get their own plans (``b200_cnn.py``,
  ``b200_resnet.py``) behind the same learner;
* the dataset is resident on the device, shuffled by one gather kernel per
  epoch, and a whole epoch is ONE CUDA graph launch (no host work per step);
* a round is a device pipeline: ``evaluate_async`` / ``fit(after_launch=...)`` /
  ``evaluate_finish`` queue evaluation, the epochs, validation and (through
  ``after_launch``) the model exchange back to back on the learner's own
  stream; the host reads the results afterwards.

Models without a fused plan (GPT-2) fall back to eager autograd for forward/backward
(``TorchLearner``) but keep the arena, so the gossip data plane is identical.
"""

from __future__ import annotations

import threading
import time
from typing import Any, Dict, List, Optional, Tuple

import torch
from torch import nn

from p2pfl_b200 import ops
from p2pfl_b200.learning.params import FlatParams
from p2pfl_b200.learning.torch_learner import TorchLearner
from p2pfl_b200.management.logger import logger
from p2pfl_b200.parallel.arena import ArenaPool, DeviceArena
from p2pfl_b200.settings import Settings

_capture_lock = threading.Lock()

# Compute streams of the learners: created by the runtime, owned by exactly one live learner at a time,
# handed to the next learner afterwards and never destroyed (tensors and events of the framework's
# allocator may still name a stream when the interpreter shuts down).
_free_streams: Dict[int, List[int]] = {}
_streams_lock = threading.Lock()


def _acquire_stream(device: int) -> int:
    with _streams_lock:
        free = _free_streams.setdefault(device, [])
        if free:
            return free.pop()
    return ops.rt_stream_create(device)


def _release_stream(device: int, stream: int) -> None:
    with _streams_lock:
        _free_streams.setdefault(device, []).append(stream)


def _pad8(x: int) -> int:
    return (x + 7) // 8 * 8


class _LinearSpec:
    def __init__(self, name: str, in_f: int, out_f: int, w_off: int, b_off: int) -> None:
        self.name, self.in_f, self.out_f, self.w_off, self.b_off = name, in_f, out_f, w_off, b_off
        self.wt_off = -1  # offset of the bf16 transposed copy, -1: not needed (first layer)


class B200Learner(TorchLearner):
    """``Node(model, data, learner=B200Learner)``; needs a CUDA device."""

    def __init__(
        self,
        model: Any,
        data: Any,
        self_addr: str = "unknown-node",
        epochs: int = 1,
        device: Optional[str] = None,
        data_mode: str = "resident",
        inbox_slots: Optional[int] = None,
    ) -> None:
        if not torch.cuda.is_available():
            raise RuntimeError("B200Learner needs a CUDA device; use TorchLearner on the host")
        ops.lib()  # the kernel library is mandatory on this path
        self.data_mode = data_mode  # "resident": upload the splits once; "stream": H2D from pinned memory every round
        self.inbox_slots = inbox_slots or Settings.DEVICE_INBOX_SLOTS
        self.arena: Optional[DeviceArena] = None
        self.fused = False
        self._layers: List[_LinearSpec] = []
        self._graphs: Dict[Tuple[str, int], Any] = {}
        self._train_dev: Optional[Tuple[torch.Tensor, torch.Tensor]] = None
        self._eval_dev: Dict[str, Tuple[torch.Tensor, torch.Tensor]] = {}
        # "stream" mode: the next round's splits travel host -> device on a copy stream while this round
        # computes (staging buffers, see _prefetch / _take_prefetched)
        self._stage: Dict[str, Dict[str, Any]] = {}
        self._copy_stream: Optional[torch.cuda.Stream] = None
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.kernel_launches_per_epoch = 0
        super().__init__(model, data, self_addr, epochs, device)
        # The compute stream captures graphs, so it must be this learner's alone: torch hands its streams
        # out of a pool of 32 per device, and with a few dozen nodes in one process two learners got the
        # same one - the work of the first landed inside the capture of the second.
        self._own_stream = _acquire_stream(self.device.index or 0)
        self.stream = torch.cuda.ExternalStream(self._own_stream, device=self.device)
        # data-plane copies issued by OTHER threads (inbound pulls) must never touch the
        # compute stream: it may be capturing a graph at that moment
        self.comm_stream = torch.cuda.Stream(device=self.device)

    # --------------------------------------------------------------- storage
    def _alloc_flat(self, total: int) -> torch.Tensor:
        if self.arena is not None:
            ArenaPool.release(self.arena)
        self.arena = ArenaPool.acquire(
            self.addr,
            total,
            slots=self.inbox_slots,
            tile=Settings.DEVICE_TILE_ELEMS,
            wire_bf16=(Settings.WIRE_DTYPE == "bf16"),
            device=self.device.index or 0,
        )
        return self.arena.params

    def set_model(self, model: Any) -> None:
        super().set_model(model)
        dev = self.device
        n = self.layout.total
        self.w16 = torch.zeros(n, dtype=torch.bfloat16, device=dev)
        self.grads = torch.zeros(n, dtype=torch.float32, device=dev)
        self.adam_m = torch.zeros(n, dtype=torch.float32, device=dev)
        self.adam_v = torch.zeros(n, dtype=torch.float32, device=dev)
        self.adam_step = torch.zeros(1, dtype=torch.int32, device=dev)
        self._graphs = {}
        self._plan_fused()
        self._params_written()

    def _plan_fused(self) -> None:
        """A model qualifies for the fully fused path when it is a plain stack of
        ``nn.Linear`` layers with ReLU in between (the reference MLP)."""
        self.fused = False
        self._epoch_kernel = False
        self._layers = []
        self.cnn = None
        self.host_seconds: Dict[str, float] = {}  # where the host thread spends a fit (perm upload / launch / sync)
        self._tdescs = []
        if getattr(self.model, "fused_spec", None) == "cnn":
            from p2pfl_b200.learning.b200_cnn import CnnPlan

            plan = CnnPlan(self)
            if plan.ok:
                self.cnn = plan
                self._feat, self._ncls = plan.feat, plan.ncls
                self._tdescs, self.wT16 = plan.tdescs, plan.wT16
                self.fused = True
            return
        if getattr(self.model, "fused_spec", None) == "resnet18":
            from p2pfl_b200.learning.b200_resnet import ResNetPlan

            plan = ResNetPlan(self)
            if plan.ok:
                self.cnn = plan  # same plan interface as the CNN (alloc / forward_eval / train_step / refresh)
                self._feat, self._ncls = plan.feat, plan.ncls
                self._tdescs, self.wT16 = plan.tdescs, plan.wT16
                self.fused = True
            return
        if getattr(self.model, "fused_spec", None) != "mlp":
            return
        linears = [(n, m) for n, m in self.model.named_children() if isinstance(m, nn.Linear)]
        if not linears or len(linears) != len(list(self.model.children())):
            return
        if int(getattr(self.data, "batch_size", 32)) % 8:
            return  # the batch is the K dimension of the wgrad GEMM: 16-byte row pitch
        prev_out = None
        for name, lin in linears:
            if lin.bias is None or (prev_out is not None and lin.in_features != prev_out):
                return
            if lin.in_features % 8:
                return  # TMA needs 16-byte row pitch on the K-major weight
            w = self.layout.by_name[f"{name}.weight"]
            b = self.layout.by_name[f"{name}.bias"]
            self._layers.append(_LinearSpec(name, lin.in_features, lin.out_features, w.offset, b.offset))
            prev_out = lin.out_features
        # transposed bf16 weights (for dgrad) of every layer but the first
        off = 0
        self._tdescs = []
        for L in self._layers[1:]:
            L.wt_off = off
            ld = _pad8(L.out_f)
            self._tdescs.append((L.w_off, off, L.out_f, L.in_f, ld))
            off += L.in_f * ld
        self.wT16 = torch.zeros(max(off, 8), dtype=torch.bfloat16, device=self.device)
        self._feat, self._ncls = self._layers[0].in_f, self._layers[-1].out_f
        self.fused = True
        # the whole-epoch cluster kernel is compiled for one architecture and one parameter layout
        dims = tuple([self._layers[0].in_f] + [L.out_f for L in self._layers])
        offs = [o for L in self._layers for o in (L.w_off, L.b_off)]
        want, o = [], 0
        for i_f, o_f in zip(ops.MLP_EPOCH_DIMS[:-1], ops.MLP_EPOCH_DIMS[1:]):
            want += [o, o + i_f * o_f]
            o += i_f * o_f + (o_f + 31) // 32 * 32
        self._epoch_kernel = (
            Settings.MLP_EPOCH_KERNEL
            and dims == ops.MLP_EPOCH_DIMS
            and offs == want
            and int(getattr(self.data, "batch_size", 32)) == ops.MLP_EPOCH_BATCH
            and ops.mlp_epoch_supported(self.device)
        )

    def _params_written(self) -> None:
        """fp32 master changed from outside: refresh the bf16 working copies."""
        if getattr(self, "w16", None) is None:
            return
        with torch.cuda.stream(self._s()):
            ops.cast_f32_bf16(self.flat, self.w16)
            if self.fused and self.cnn is not None:
                self.cnn.refresh()
            elif self.fused and self._tdescs:
                ops.transpose_cast(self.flat, self.wT16, self._tdescs)

    def _params_written_by_plane(self) -> None:
        """The data plane wrote fp32 params AND the bf16 copy; only the transposes are stale."""
        if self.fused and (self._tdescs or self.cnn is not None):
            with torch.cuda.stream(self._s()):
                if self.cnn is not None:
                    self.cnn.refresh()
                else:
                    ops.transpose_cast(self.flat, self.wT16, self._tdescs)

    def exchange_snapshot(self) -> Optional[torch.Tensor]:
        """fp32 buffer the fused exchange copies the pre-exchange model into (tile by tile, as it
        pushes), so that an exchange that does not complete can be undone."""
        if not Settings.FUSED_SNAPSHOT:
            return None
        snap = getattr(self, "_snapshot", None)
        if snap is None or snap.numel() != self.flat.numel():
            snap = self._snapshot = torch.empty_like(self.flat)
        return snap

    def _s(self) -> torch.cuda.Stream:
        return getattr(self, "stream", None) or torch.cuda.current_stream(self.device)

    def close(self) -> None:
        self._graphs = {}
        self._sync_streams()
        if self.arena is not None:
            # views into the arena must not outlive it
            self.flat = None
            ArenaPool.release(self.arena)
            self.arena = None
        own, self._own_stream = getattr(self, "_own_stream", None), None
        if own:
            _release_stream(self.device.index or 0, own)  # the next learner of this process may take it

    def _sync_streams(self) -> None:
        """Wait for this learner's own streams (never ``torch.cuda.synchronize``: CUDA refuses a device-wide
        synchronize while another node of the process captures a graph)."""
        for name in ("stream", "comm_stream", "_copy_stream"):
            st = getattr(self, name, None)
            if st is not None:
                st.synchronize()

    # ------------------------------------------------------------ parameters
    def export_model(self) -> FlatParams:
        with torch.cuda.stream(self._s()):
            out = FlatParams(self.flat.detach().clone(), self.layout, self._extra())
        self._s().synchronize()
        return out

    def export_payload(self, params: Any = None) -> Any:
        """Device payload for data planes that move device memory themselves: the
        model stays on the GPU; what travels on the control plane is a reference."""
        from p2pfl_b200.parallel.payload import DevicePayload

        if params is None:
            params = self.export_model()
        if not isinstance(params, FlatParams):
            raise TypeError("export_payload needs a FlatParams")
        return DevicePayload(params, self)

    def decode_parameters(self, data: Any) -> FlatParams:
        """Host payloads decode as usual; a device reference is resolved by pulling the
        tensor from the sender's arena over NVLink."""
        from p2pfl_b200.parallel.payload import DeviceRef

        if DeviceRef.is_ref(data):
            plane = getattr(self, "plane", None)
            if plane is None:
                from p2pfl_b200.exceptions import DecodingParamsError

                raise DecodingParamsError("device reference received but no device data plane is attached")
            return plane.pull(DeviceRef.from_bytes(data))
        return super().decode_parameters(data)

    def live_params(self) -> FlatParams:
        """The live model WITHOUT a copy (valid until the next update)."""
        return FlatParams(self.flat, self.layout, self._extra())

    def set_parameters(self, params: Any) -> None:
        if isinstance(params, FlatParams) and params.flat.data_ptr() == self.flat.data_ptr():
            self._params_written()  # already in place (device data plane wrote it)
            self._s().synchronize()
            return
        with torch.cuda.stream(self._s()):
            if isinstance(params, FlatParams) and params.flat.is_cuda:
                # the producer may have written it on another stream
                self._s().wait_stream(torch.cuda.current_stream(params.flat.device))
            super().set_parameters(params)
        self._s().synchronize()

    # ------------------------------------------------------------------ data
    def _upload(self, split: str) -> Tuple[torch.Tensor, torch.Tensor]:
        x, y = self.data.pinned_tensors(split) if hasattr(self.data, "pinned_tensors") else self.data.tensors(split)
        xd = self._rows(x.to(self.device, non_blocking=True))
        yd = y.to(self.device, non_blocking=True).to(torch.int32)  # int64 travels, the cast is on the device
        self.h2d_bytes += x.numel() * x.element_size() + y.numel() * y.element_size()
        return xd, yd

    def _rows(self, x_dev: torch.Tensor) -> torch.Tensor:
        """[n, ...] device tensor -> [n, feat] rows; channels-last for plans that consume NHWC."""
        if x_dev.dim() == 4 and x_dev.shape[1] > 1 and getattr(self.cnn, "nhwc_input", False):
            x_dev = x_dev.permute(0, 2, 3, 1)
        return x_dev.reshape(x_dev.shape[0], -1).contiguous()

    def _prefetch(self, split: str) -> None:
        """Stream mode: start the host -> device copy of ``split`` for the NEXT round now, on the
        copy stream, into staging buffers (the live buffers are still being read by this round)."""
        if self.data_mode != "stream" or not hasattr(self.data, "pinned_tensors"):
            return
        dst = self._train_dev if split == "train" else self._eval_dev.get(split)
        if dst is None:
            return
        x, y = self.data.pinned_tensors(split)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        cs = self._copy_stream
        st = self._stage.get(split)
        if st is None:
            st = self._stage[split] = {
                "x": torch.empty(x.shape, dtype=x.dtype, device=self.device),
                "y": torch.empty(y.shape, dtype=y.dtype, device=self.device),
                "ready": None, "taken": None,
            }  # fmt: skip
            for t in (st["x"], st["y"]):
                t.record_stream(cs)  # written on the copy stream: the allocator must not recycle it early
        if st["ready"] is not None:
            return  # one already in flight and not consumed yet
        if st["taken"] is not None:
            cs.wait_event(st["taken"])  # the last consumer's device copy out of the staging buffer
        with torch.cuda.stream(cs):
            st["x"].copy_(x, non_blocking=True)
            st["y"].copy_(y, non_blocking=True)
            ready = torch.cuda.Event()
            ready.record(cs)
        st["ready"] = ready
        self.h2d_bytes += x.numel() * x.element_size() + y.numel() * y.element_size()

    def _take_prefetched(self, split: str, dst: Tuple[torch.Tensor, torch.Tensor]) -> bool:
        """Move a finished prefetch into the live buffers (device copy on the learner's stream)."""
        st = self._stage.get(split)
        if st is None or st["ready"] is None:
            return False
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(st["ready"])
        dst[0].copy_(self._rows(st["x"]))
        dst[1].copy_(st["y"])  # int64 travels, the cast is on the device
        taken = torch.cuda.Event()
        taken.record(cur)
        st["ready"], st["taken"] = None, taken
        return True

    def _train_data(self) -> Tuple[torch.Tensor, torch.Tensor]:
        if self.data_mode == "stream" or self._train_dev is None:
            if self._train_dev is not None and self.data_mode == "stream":
                # same device buffers every round so captured graphs stay valid
                if not self._take_prefetched("train", self._train_dev):
                    x, y = self.data.pinned_tensors("train")
                    self._train_dev[0].copy_(self._rows(x.to(self.device, non_blocking=True)))
                    self._train_dev[1].copy_(y.to(self.device, non_blocking=True))
                    self.h2d_bytes += x.numel() * x.element_size() + y.numel() * y.element_size()
            else:
                self._train_dev = self._upload("train")
        return self._train_dev

    def _eval_data(self, split: str) -> Tuple[torch.Tensor, torch.Tensor]:
        cur = self._eval_dev.get(split)
        if cur is None:
            self._eval_dev[split] = cur = self._upload(split)
        elif self.data_mode == "stream":
            if not self._take_prefetched(split, cur):
                x, y = self.data.pinned_tensors(split)
                cur[0].copy_(self._rows(x.to(self.device, non_blocking=True)))
                cur[1].copy_(y.to(self.device, non_blocking=True))
                self.h2d_bytes += x.numel() * x.element_size() + y.numel() * y.element_size()
        return cur

    # --------------------------------------------------------- fused forward
    def _forward(self, x16: torch.Tensor, ldx: int, bsz: int, act: Dict[str, Any], train: bool) -> None:
        """Swap-AB forward through every layer: D[out, bsz] = W[out, in] . X[bsz, in]^T."""
        Ls = self._layers
        for li, L in enumerate(Ls):
            last = li == len(Ls) - 1
            A = self.w16[L.w_off :]
            Bm, ldb = (x16, ldx) if li == 0 else (act["H"][li - 1], _pad8(L.in_f))
            bias = self.flat[L.b_off :]
            if last:
                ops.gemm_tn(A, L.in_f, Bm, ldb, L.out_f, bsz, L.in_f, bias=bias, bias_per_row=True,
                            out_f32_t=act["logits"], ld_f32_t=act["ld_logits"])  # fmt: skip
            else:
                ops.gemm_tn(A, L.in_f, Bm, ldb, L.out_f, bsz, L.in_f, bias=bias, bias_per_row=True, relu=True,
                            out_bf16=act["HT"][li] if train else None, ld_bf16=act["ldB"],
                            out_bf16_t=act["H"][li], ld_bf16_t=_pad8(L.out_f))  # fmt: skip

    def _alloc_act(self, B: int, train: bool) -> Dict[str, Any]:
        if self.cnn is not None:
            return self.cnn.alloc(B, train)
        dev = self.device
        Ls = self._layers
        ldB = _pad8(B)
        C = Ls[-1].out_f
        act: Dict[str, Any] = {"ldB": ldB, "ld_logits": _pad8(C)}
        act["H"] = [torch.zeros(B, _pad8(L.out_f), dtype=torch.bfloat16, device=dev) for L in Ls[:-1]]
        act["logits"] = torch.zeros(B, _pad8(C), dtype=torch.float32, device=dev)
        if train:
            act["HT"] = [torch.zeros(L.out_f, ldB, dtype=torch.bfloat16, device=dev) for L in Ls[:-1]]
            act["dZT"] = [torch.zeros(L.out_f, ldB, dtype=torch.bfloat16, device=dev) for L in Ls]
            act["dZ"] = [torch.zeros(B, _pad8(L.out_f), dtype=torch.bfloat16, device=dev) for L in Ls]
        return act

    def _train_step(self, x16, xT16, y, bsz: int, act, loss_slot, step_offset: int, lr: float) -> None:
        if self.cnn is not None:
            self.cnn.train_step(x16, y, bsz, act, loss_slot, step_offset, lr)
            return
        Ls = self._layers
        nL = len(Ls)
        feat = Ls[0].in_f
        ldB = act["ldB"]
        C = Ls[-1].out_f
        if self._epoch_kernel and bsz <= ops.MLP_EPOCH_BATCH:
            # one step through the epoch kernel (tests / odd call sites; an epoch is ONE such call)
            if step_offset != 1:
                raise RuntimeError("single steps through the epoch kernel start at step offset 1")
            ops.mlp_epoch(x16, y, bsz, self.flat, self.adam_m, self.adam_v, self.w16, loss_slot, self.adam_step, lr,
                          grads=self.grads)  # fmt: skip
            return
        if Settings.MLP_CHAIN_CLUSTER > 0 and 3 * nL - 1 <= 8 and C <= 256 and bsz <= 32:
            cluster = min(Settings.MLP_CHAIN_CLUSTER, ops.gemm_chain_max_cluster(self.device))
            self._train_step_chain(x16, xT16, y, bsz, act, loss_slot, step_offset, lr, cluster=cluster)
            return
        if Settings.MLP_CHAIN and 2 * nL - 1 <= 8 and C <= 256:
            self._train_step_chain(x16, xT16, y, bsz, act, loss_slot, step_offset, lr)
            return
        # no gradient memset: every weight gradient is overwritten by its wgrad GEMM and every
        # bias gradient is stored (not accumulated) by the CE kernel / the dgrad row sums
        # (batches wider than one 32-column tile accumulate row sums, so they do need zeros)
        if bsz > 32:
            self.grads.zero_()
        self._forward(x16, feat, bsz, act, train=True)
        ops.softmax_xent(act["logits"], act["ld_logits"], y, bsz, C, act["dZ"][-1], _pad8(C), act["dZT"][-1], ldB,
                         self.grads[Ls[-1].b_off :], loss_slot)  # fmt: skip
        for li in range(nL - 1, -1, -1):
            L = Ls[li]
            # wgrad: dW[out, in] = dZ^T[out, bsz] . (X^T[in, bsz])^T   (K = batch)
            Bm = xT16 if li == 0 else act["HT"][li - 1]
            group = [ops.gemm_args(act["dZT"][li], ldB, Bm, ldB, L.out_f, L.in_f, bsz,
                                   out_f32=self.grads[L.w_off :], ld_f32=L.in_f)]  # fmt: skip
            if li > 0:
                # dgrad: dH^T[in, bsz] = W^T[in, out] . dZ[bsz, out]^T, masked by ReLU, bias grad = row sums
                P = Ls[li - 1]
                group.append(ops.gemm_args(
                    self.wT16[L.wt_off :], _pad8(L.out_f), act["dZ"][li], _pad8(L.out_f), L.in_f, bsz, L.out_f,
                    mask=act["HT"][li - 1], ld_mask=ldB, out_bf16=act["dZT"][li - 1], ld_bf16=ldB,
                    out_bf16_t=act["dZ"][li - 1] if li - 1 > 0 else None, ld_bf16_t=_pad8(P.out_f),
                    rowsum=self.grads[P.b_off :]))  # fmt: skip
            # a layer's wgrad and dgrad both depend only on its dZ: one launch for the pair
            ops.gemm_tn_grouped(self.device, group)
        # Adam over the whole arena, refreshing the bf16 copy AND the transposed dgrad operands
        ops.adam_step_t(self.flat, self.grads, self.adam_m, self.adam_v, self.w16, lr, self.adam_step, step_offset,
                        self.wT16, self._tdescs)  # fmt: skip

    def _train_step_chain(self, x16, xT16, y, bsz: int, act, loss_slot, step_offset: int, lr: float, cluster: int = 0) -> None:
        """``cluster`` = 0, three launches per step: (1) ONE single-CTA kernel for every dependent
        link - the forward chain, the loss and the dgrad chain; (2) all weight-gradient GEMMs, which
        depend only on (1), as one grouped launch; (3) Adam + bf16 refresh + transposes.
        ``cluster`` > 0, two launches: (1) a thread-block cluster walks the same links and then all
        weight gradients as one more link; (2) Adam."""
        Ls = self._layers
        nL = len(Ls)
        feat = Ls[0].in_f
        ldB = act["ldB"]
        C = Ls[-1].out_f
        if bsz > 32:
            self.grads.zero_()  # row sums of wider batches are accumulated
        gemms, links = [], []
        for li, L in enumerate(Ls):  # forward, swap-AB: D[out, bsz] = W[out, in] . X[bsz, in]^T
            last = li == nL - 1
            Bm, ldb = (x16, feat) if li == 0 else (act["H"][li - 1], _pad8(L.in_f))
            common = dict(bias=self.flat[L.b_off :], bias_per_row=True)
            if cluster > 0:
                # batch-major (D[bsz, out] = X . W^T): the N tiles of a link spread over the cluster's CTAs,
                # and each streams a 32-row slice of W instead of 128 rows - the forward links are bound by
                # what ONE SM can pull through TMA
                colb = dict(bias=self.flat[L.b_off :], bias_per_row=False)
                if last:
                    gemms.append(ops.gemm_args(Bm, ldb, self.w16[L.w_off :], L.in_f, bsz, L.out_f, L.in_f,
                                               out_f32=act["logits"], ld_f32=act["ld_logits"], **colb))  # fmt: skip
                else:
                    gemms.append(ops.gemm_args(Bm, ldb, self.w16[L.w_off :], L.in_f, bsz, L.out_f, L.in_f, relu=True,
                                               out_bf16=act["H"][li], ld_bf16=_pad8(L.out_f), out_bf16_t=act["HT"][li],
                                               ld_bf16_t=ldB, **colb))  # fmt: skip
            elif last:
                gemms.append(ops.gemm_args(self.w16[L.w_off :], L.in_f, Bm, ldb, L.out_f, bsz, L.in_f,
                                           out_f32_t=act["logits"], ld_f32_t=act["ld_logits"], **common))  # fmt: skip
            else:
                gemms.append(ops.gemm_args(self.w16[L.w_off :], L.in_f, Bm, ldb, L.out_f, bsz, L.in_f, relu=True,
                                           out_bf16=act["HT"][li], ld_bf16=ldB, out_bf16_t=act["H"][li],
                                           ld_bf16_t=_pad8(L.out_f), **common))  # fmt: skip
            links.append(("gemm", len(gemms) - 1))
        links.append(("ce", 0))
        ce = ops.ce_args(act["logits"], act["ld_logits"], y, bsz, C, act["dZ"][-1], _pad8(C), act["dZT"][-1], ldB,
                         self.grads[Ls[-1].b_off :], loss_slot)  # fmt: skip
        for li in range(nL - 1, 0, -1):  # dgrad: dH^T[in, bsz] = W^T[in, out] . dZ[bsz, out]^T, ReLU mask, bias grad
            L, P = Ls[li], Ls[li - 1]
            gemms.append(ops.gemm_args(
                self.wT16[L.wt_off :], _pad8(L.out_f), act["dZ"][li], _pad8(L.out_f), L.in_f, bsz, L.out_f,
                mask=act["HT"][li - 1], ld_mask=ldB, out_bf16=act["dZT"][li - 1], ld_bf16=ldB,
                out_bf16_t=act["dZ"][li - 1] if li - 1 > 0 else None, ld_bf16_t=_pad8(P.out_f),
                rowsum=self.grads[P.b_off :]))  # fmt: skip
            links.append(("gemm", len(gemms) - 1))
        wg = []
        for li, L in enumerate(Ls):  # wgrad: dW[out, in] = dZ^T[out, bsz] . (X^T[in, bsz])^T   (K = batch)
            Bm = xT16 if li == 0 else act["HT"][li - 1]
            wg.append(ops.gemm_args(act["dZT"][li], ldB, Bm, ldB, L.out_f, L.in_f, bsz,
                                    out_f32=self.grads[L.w_off :], ld_f32=L.in_f))  # fmt: skip
        if cluster > 0:
            links.append(("group", len(gemms), len(wg)))
            ops.gemm_chain(self.device, gemms + wg, links, ce, cluster=cluster)
        else:
            ops.gemm_chain(self.device, gemms, links, ce)
            for i in range(0, len(wg), 4):
                ops.gemm_tn_grouped(self.device, wg[i : i + 4])
        ops.adam_step_t(self.flat, self.grads, self.adam_m, self.adam_v, self.w16, lr, self.adam_step, step_offset,
                        self.wT16, self._tdescs)  # fmt: skip

    def debug_single_step(self, x: torch.Tensor, y: torch.Tensor) -> float:
        """One eager (un-graphed) fused step on the given batch; returns the mean loss.
        Test hook: leaves the gradients in ``self.grads``."""
        B = x.shape[0]
        feat = self._feat
        with torch.cuda.stream(self._s()):
            xf = self._rows(x.to(self.device, torch.float32))
            yi = y.to(self.device).to(torch.int32)
            xs = torch.zeros(B, feat, dtype=torch.bfloat16, device=self.device)
            xsT = torch.zeros(1, feat, _pad8(B), dtype=torch.bfloat16, device=self.device)
            ys = torch.zeros(B, dtype=torch.int32, device=self.device)
            ops.gather_rows(xf, None, B, feat, _pad8(B), xs, xsT, yi, ys)
            act = self._alloc_act(B, train=True)
            loss = torch.zeros(1, dtype=torch.float32, device=self.device)
            self._train_step(xs, xsT[0], ys, B, act, loss, 1, float(getattr(self.model, "lr", 1e-3)))
            ops.add_i32(self.adam_step, 1)
        self._s().synchronize()
        return float(loss) / B

    # ------------------------------------------------------------ epoch graph
    def _build_epoch(self, n: int, B: int) -> Dict[str, Any]:
        """Capture one whole epoch (gather + every step) as a single CUDA graph."""
        dev = self.device
        feat = self._feat
        nb = (n + B - 1) // B
        x, y = self._train_dev
        ep: Dict[str, Any] = {"n": n, "B": B, "nb": nb}
        ep["perm"] = torch.arange(n, dtype=torch.int32, device=dev)
        ep["perm_host"] = torch.empty(n, dtype=torch.int32).pin_memory()
        ep["perm_host2"] = torch.empty(n, dtype=torch.int32).pin_memory()
        ep["xs"] = torch.zeros(nb * B, feat, dtype=torch.bfloat16, device=dev)
        # the MLP's first wgrad GEMM wants each batch transposed as well; the CNN does not
        need_t = self.cnn is None
        ep["xsT"] = torch.zeros(nb if need_t else 1, feat, _pad8(B), dtype=torch.bfloat16, device=dev)
        ep["ys"] = torch.zeros(nb * B, dtype=torch.int32, device=dev)
        ep["loss"] = torch.zeros(nb, dtype=torch.float32, device=dev)
        ep["act"] = self._alloc_act(B, train=True)
        lr = float(getattr(self.model, "lr", 1e-3))

        def body() -> None:
            if self._epoch_kernel:
                # shuffle, then the whole epoch in ONE cluster launch (weights resident on chip)
                ops.gather_rows(x, ep["perm"], n, feat, _pad8(B), ep["xs"], None, y, ep["ys"])
                ops.mlp_epoch(ep["xs"], ep["ys"], n, self.flat, self.adam_m, self.adam_v, self.w16, ep["loss"],
                              self.adam_step, lr)  # fmt: skip
                ops.add_i32(self.adam_step, nb)
                return
            ep["loss"].zero_()
            ops.gather_rows(x, ep["perm"], n, feat, _pad8(B), ep["xs"], ep["xsT"] if need_t else None, y, ep["ys"])
            for i in range(nb):
                bsz = min(B, n - i * B)
                self._train_step(ep["xs"][i * B :], ep["xsT"][i if need_t else 0], ep["ys"][i * B :], bsz, ep["act"],
                                 ep["loss"][i:], i + 1, lr)  # fmt: skip
            ops.add_i32(self.adam_step, nb)

        # warm-up outside capture (first-use attribute calls, lazy driver entry points)
        before = ops.launch_count()
        saved = [t.clone() for t in (self.flat, self.adam_m, self.adam_v, self.adam_step, self.w16, self.wT16)]
        body()
        for t, s in zip((self.flat, self.adam_m, self.adam_v, self.adam_step, self.w16, self.wT16), saved):
            t.copy_(s)
        ep["launches"] = ops.launch_count() - before
        self._s().synchronize()
        g = torch.cuda.CUDAGraph()
        with _capture_lock:
            with torch.cuda.graph(g, stream=self._s(), capture_error_mode="thread_local"):
                body()
        ep["graph"] = g
        return ep

    def fit(self, after_launch: Optional[Any] = None) -> None:
        """``after_launch`` (fused plans only): called on this thread once the LAST epoch, its
        validation pass and the copies of their results are queued on the stream, before anything
        waits for them - whatever it queues (the round's model exchange) starts the moment training
        ends on the device, with no host round trip in between."""
        if self.epochs <= 0 or self.model is None:
            return
        if not self.fused:
            with torch.cuda.stream(self._s()):
                super().fit()
                self._params_written()
            self._s().synchronize()
            return
        self._interrupt.clear()
        B = int(getattr(self.data, "batch_size", 32))
        with torch.cuda.stream(self._s()):
            x, _ = self._train_data()
            n = x.shape[0]
            key = ("epoch", n * 100003 + B)
            if key not in self._graphs:
                self._graphs[key] = self._build_epoch(n, B)
            ep = self._graphs[key]
            if Settings.RESET_OPTIMIZER_EACH_ROUND:
                # the reference builds a fresh Trainer (=> fresh Adam) per round (Q16)
                self.adam_m.zero_()
                self.adam_v.zero_()
                self.adam_step.zero_()
            loss_ready = None
            for e in range(self.epochs):
                if self._interrupt.is_set():
                    logger.info(self.addr, "fit interrupted")
                    break
                # host RNG: the CUDA generator cannot be used while another node's
                # thread captures a graph on this device
                ta = time.perf_counter()
                if not ep.get("perm_ready"):
                    ep["perm_host"].copy_(torch.randperm(n, dtype=torch.int32))
                ep["perm"].copy_(ep["perm_host"], non_blocking=True)
                up = torch.cuda.Event()
                up.record()
                self.h2d_bytes += n * 4
                tb = time.perf_counter()
                self._prefetch("val")  # stream mode: this epoch's validation split travels during the epoch
                ep["graph"].replay()
                ops.add_launches(ep["launches"])
                self.kernel_launches_per_epoch = ep["launches"]
                # the next epoch's shuffle is drawn while the device runs this one (the upload above
                # is done with the first pinned buffer before anything waits on the stream again)
                ep["perm_host"], ep["perm_host2"] = ep["perm_host2"], ep["perm_host"]
                if ep.get("perm_up2") is not None:
                    ep["perm_up2"].synchronize()  # the upload that last read this buffer (long done)
                ep["perm_up2"] = up
                ep["perm_host"].copy_(torch.randperm(n, dtype=torch.int32))
                ep["perm_ready"] = True
                if e == self.epochs - 1:
                    # stream mode: next round's splits start travelling while the device runs this epoch
                    self._prefetch("train")
                    self._prefetch("test")
                tc = time.perf_counter()
                # per-epoch validation, as Lightning does for the reference (mlp.py:82-91)
                last = e == self.epochs - 1
                vh = self._eval_launch("val")
                if last:
                    loss_host = ep.get("loss_host")
                    if loss_host is None:
                        loss_host = ep["loss_host"] = torch.empty(ep["nb"], dtype=torch.float32).pin_memory()
                    loss_host.copy_(ep["loss"], non_blocking=True)
                    loss_ready = torch.cuda.Event()
                    loss_ready.record()
                    if after_launch is not None:
                        after_launch()
                val = self._eval_result(vh)
                td = time.perf_counter()
                ht = self.host_seconds
                ht["fit_perm"] = ht.get("fit_perm", 0.0) + tb - ta
                ht["fit_replay_call"] = ht.get("fit_replay_call", 0.0) + tc - tb
                ht["fit_val_sync"] = ht.get("fit_val_sync", 0.0) + td - tc
                if val is not None:
                    logger.log_metric(self.addr, "val_loss", val[0], step=(e + 1) * ep["nb"])
                    logger.log_metric(self.addr, "val_metric", val[1], step=(e + 1) * ep["nb"])
            if loss_ready is not None:
                loss_ready.synchronize()
                losses = ep["loss_host"].clone()  # one D2H per fit, queued before the exchange
            else:
                losses = ep["loss"].cpu()
            self.d2h_bytes += losses.numel() * 4
        sizes = [min(B, n - i * B) for i in range(ep["nb"])]
        for i in range(0, ep["nb"], self.LOG_EVERY):
            logger.log_metric(self.addr, "train_loss", float(losses[i]) / sizes[i], step=i)
        self._round_steps = ep["nb"] * self.epochs

    # -------------------------------------------------------------- evaluate
    def _build_eval(self, split: str, n: int) -> Dict[str, Any]:
        dev = self.device
        feat, C = self._feat, self._ncls
        x, y = self._eval_dev[split]
        ev: Dict[str, Any] = {"n": n}
        ev["x16"] = torch.zeros(n, feat, dtype=torch.bfloat16, device=dev)
        if self.cnn is not None:
            ev["act"] = self._alloc_act(min(n, self.cnn.EVAL_CHUNK), train=False)
            ev["logits"] = torch.zeros(n, _pad8(C), dtype=torch.float32, device=dev)
        else:
            ev["act"] = self._alloc_act(n, train=False)
            ev["logits"] = ev["act"]["logits"]
        ev["loss"] = torch.zeros(1, dtype=torch.float32, device=dev)
        ev["correct"] = torch.zeros(1, dtype=torch.int32, device=dev)
        ev["out"] = torch.zeros(2, dtype=torch.float32, device=dev)

        def body() -> None:
            ev["loss"].zero_()
            ev["correct"].zero_()
            ops.gather_rows(x, None, n, feat, 8, ev["x16"])
            if self.cnn is not None:
                self.cnn.forward_eval(ev["x16"], n, ev["act"], ev["logits"])
            else:
                self._forward(ev["x16"], feat, n, ev["act"], train=False)
            ops.eval_reduce(ev["logits"], ev["act"]["ld_logits"], y, n, C, ev["loss"], ev["correct"])
            ev["out"][0].copy_(ev["loss"][0])
            ev["out"][1].copy_(ev["correct"][0])

        before = ops.launch_count()
        body()
        ev["launches"] = ops.launch_count() - before
        self._s().synchronize()
        g = torch.cuda.CUDAGraph()
        with _capture_lock:
            with torch.cuda.graph(g, stream=self._s(), capture_error_mode="thread_local"):
                body()
        ev["graph"] = g
        return ev

    def _eval_launch(self, split: str) -> Optional[Tuple[Any, torch.Tensor, Any, int]]:
        """Queue the evaluation of a split (one graph launch) and the 8-byte copy of its result into
        pinned memory; nothing waits.  ``_eval_result`` collects."""
        x, _ = self._eval_data(split)
        n = x.shape[0]
        if n == 0:
            return None
        key = ("eval:" + split, n)
        if key not in self._graphs:
            self._graphs[key] = self._build_eval(split, n)
        ev = self._graphs[key]
        ev["graph"].replay()
        ops.add_launches(ev["launches"])
        if "out_host" not in ev:
            ev["out_host"] = torch.empty(2, dtype=torch.float32).pin_memory()
        ev["out_host"].copy_(ev["out"], non_blocking=True)
        ready = torch.cuda.Event()
        ready.record()
        self.d2h_bytes += 8
        return ev, ev["out_host"], ready, n

    def _eval_result(self, handle: Optional[Tuple[Any, torch.Tensor, Any, int]]) -> Optional[Tuple[float, float]]:
        if handle is None:
            return None
        _, host, ready, n = handle
        ready.synchronize()
        return float(host[0]) / n, float(host[1]) / n

    def _eval_split(self, split: str) -> Optional[Tuple[float, float]]:
        """(mean loss, accuracy) of the live model on a split: one graph launch, one 8-byte D2H."""
        return self._eval_result(self._eval_launch(split))

    def evaluate_async(self) -> Any:
        """Queue the test evaluation and return a handle for ``evaluate_finish`` (fused plans only;
        None otherwise: call ``evaluate``)."""
        if self.epochs <= 0 or self.model is None or not self.fused:
            return None
        with torch.cuda.stream(self._s()):
            return self._eval_launch("test")

    def evaluate_finish(self, handle: Any) -> Dict[str, float]:
        res = self._eval_result(handle)
        if res is None:
            return {}
        out = {"test_loss": res[0], "test_metric": res[1]}
        for k, v in out.items():
            logger.log_metric(self.addr, k, v)
        return out

    def evaluate(self) -> Dict[str, float]:
        if self.epochs <= 0 or self.model is None:
            return {}
        if not self.fused:
            with torch.cuda.stream(self._s()):
                return super().evaluate()
        with torch.cuda.stream(self._s()):
            res = self._eval_split("test")
        if res is None:
            return {}
        out = {"test_loss": res[0], "test_metric": res[1]}
        for k, v in out.items():
            logger.log_metric(self.addr, k, v)
        return out
