"""Generic eager-PyTorch learner for any ``nn.Module`` classifier.

Plays the role of the reference's ``LightningLearner``
(p2pfl/learning/pytorch/lightning_learner.py:56-236) without Lightning:
cross-entropy loss, Adam(lr=model.lr), ``epochs`` passes over the train split
per ``fit``, accuracy/loss on the test split for ``evaluate``.  It is the
fallback for models that have no fused sm_100a learner and the CPU oracle the
fused learner is tested against.

Parameters live in one flat fp32 vector (``ParamLayout``); the module's
parameters and float buffers are re-pointed to views of it, so
``export_model`` / ``set_parameters`` are a single copy and FedAvg runs on
whole vectors.
"""

from __future__ import annotations

import threading
from collections import OrderedDict
from typing import Any, Dict, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from p2pfl_b200.exceptions import DecodingParamsError, ModelNotMatchingError
from p2pfl_b200.learning import codec
from p2pfl_b200.learning.learner import NodeLearner
from p2pfl_b200.learning.params import FlatParams, ParamLayout
from p2pfl_b200.management.logger import logger
from p2pfl_b200.settings import Settings

_staging_lock = threading.Lock()

torch.set_num_threads(max(1, min(4, torch.get_num_threads())))


def resolve_device(device: Optional[str]) -> torch.device:
    if device in (None, "auto"):
        return torch.device(f"cuda:{torch.cuda.current_device()}") if torch.cuda.is_available() else torch.device("cpu")
    return torch.device(device)


class TorchLearner(NodeLearner):
    LOG_EVERY = 50

    def __init__(
        self,
        model: Any,
        data: Any,
        self_addr: str = "unknown-node",
        epochs: int = 1,
        device: Optional[str] = None,
    ) -> None:
        self.addr = self_addr
        self.epochs = epochs
        self.device = resolve_device(device)
        self.model: Any = None
        self.data: Any = None
        self.layout: Optional[ParamLayout] = None
        self.flat: Optional[torch.Tensor] = None
        self._optimizer: Optional[torch.optim.Optimizer] = None
        self._interrupt = threading.Event()
        self._round_steps = 0
        # bytes moved between host and device by this learner (bench.py's end-to-end accounting)
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        if model is not None:
            self.set_model(model)
        self.set_data(data)

    # ------------------------------------------------------------ model/data
    def set_model(self, model: Any) -> None:
        self.model = model.to(self.device)
        sd = self.model.state_dict()
        self.layout = ParamLayout.from_state_dict(sd)
        self.flat = self._alloc_flat(self.layout.total)
        views = self.layout.views(self.flat)
        with torch.no_grad():
            for name, t in sd.items():
                slot = self.layout.by_name[name]
                if slot.in_arena:
                    views[name].copy_(t.to(torch.float32))
        # re-point parameters/buffers to the arena (state_dict() returns detached
        # aliases, so assign through the owning modules)
        for name, p in self.model.named_parameters():
            if p.dtype == torch.float32:
                p.data = views[name]
        for name, b in self.model.named_buffers():
            if name in views and b.dtype == torch.float32:
                b.data = views[name]
        self._optimizer = None

    def _alloc_flat(self, total: int) -> torch.Tensor:
        """Storage of the flat parameter vector (subclasses place it in a device arena)."""
        return torch.zeros(total, dtype=torch.float32, device=self.device)

    def _params_written(self) -> None:
        """Hook: the flat vector was overwritten from outside (set_parameters / aggregation)."""

    def set_data(self, data: Any) -> None:
        self.data = data

    def set_epochs(self, epochs: int) -> None:
        self.epochs = epochs

    # ------------------------------------------------------------ parameters
    def _extra(self) -> Dict[str, torch.Tensor]:
        assert self.layout is not None
        sd = self.model.state_dict()
        return {s.name: sd[s.name].detach().clone() for s in self.layout.slots if not s.in_arena}

    def get_parameters(self) -> "OrderedDict[str, torch.Tensor]":
        return OrderedDict((k, v.detach()) for k, v in self.model.state_dict().items())

    def export_model(self) -> FlatParams:
        """Snapshot of the live model as a flat vector (one device copy)."""
        assert self.flat is not None and self.layout is not None
        return FlatParams(self.flat.detach().clone(), self.layout, self._extra())

    def _as_state_dict(self, params: Any) -> "OrderedDict[str, torch.Tensor]":
        if params is None:
            return self.get_parameters()
        if isinstance(params, FlatParams):
            return params.state_dict()
        return params

    def encode_parameters(self, params: Any = None) -> bytes:
        """Host wire format (reference: lightning_learner.py:113-122).  A flat model on the device
        crosses to the host as ONE copy into a pinned staging buffer, not one
        ``.to("cpu")`` per tensor; the per-entry arrays are views of that buffer."""
        assert self.layout is not None
        flat_src: Optional[torch.Tensor] = None
        extra: Dict[str, torch.Tensor] = {}
        if params is None and self.flat is not None:
            flat_src, extra = self.flat.detach(), self._extra()
        elif isinstance(params, FlatParams) and params.layout == self.layout:
            flat_src, extra = params.flat.detach(), params.extra
        if flat_src is not None and flat_src.is_cuda:
            with _staging_lock:  # the gossiper and the learning thread may both encode
                host = self._host_staging(flat_src.numel())
                host.copy_(flat_src.to(torch.float32), non_blocking=True)
                torch.cuda.current_stream(flat_src.device).synchronize()
                views = self.layout.views(host)
                arrays = []
                for s in self.layout.slots:
                    t = views[s.name] if s.in_arena else torch.as_tensor(extra[s.name]).detach().cpu()
                    arrays.append(t.to(s.dtype).numpy())
                return codec.encode_arrays(arrays)  # pickled (copied) before the staging buffer is reused
        sd = self._as_state_dict(params)
        arrays = []
        for s in self.layout.slots:
            t = torch.as_tensor(sd[s.name]).detach()
            arrays.append(t.to("cpu", s.dtype).numpy())
        return codec.encode_arrays(arrays)

    def _host_staging(self, numel: int) -> torch.Tensor:
        """Pinned fp32 buffer reused by every device -> host encode of this learner."""
        buf = getattr(self, "_staging", None)
        if buf is None or buf.numel() != numel:
            buf = torch.empty(numel, dtype=torch.float32)
            try:
                buf = buf.pin_memory()
            except Exception:
                pass
            self._staging = buf
        return buf

    def decode_parameters(self, data: Any) -> FlatParams:
        """bytes -> FlatParams on this learner's device.  Count, shapes and
        dtypes are validated here (Q20)."""
        assert self.layout is not None
        if isinstance(data, FlatParams):
            self.check_parameters(data)
            if data.flat.device != self.device:
                return FlatParams(data.flat.to(self.device), data.layout, data.extra)
            return data
        if not isinstance(data, (bytes, bytearray, memoryview)):
            raise DecodingParamsError(f"unsupported payload type {type(data).__name__}")
        arrays = codec.decode_arrays(bytes(data))
        slots = self.layout.slots
        if len(arrays) != len(slots):
            raise ModelNotMatchingError(f"payload has {len(arrays)} tensors, local model has {len(slots)}")
        host = np.zeros(self.layout.total, dtype=np.float32)
        extra: Dict[str, torch.Tensor] = {}
        for s, a in zip(slots, arrays):
            if tuple(a.shape) != s.shape:
                raise ModelNotMatchingError(f"{s.name}: payload shape {tuple(a.shape)} != local {s.shape}")
            if s.in_arena:
                if a.dtype.kind != "f":
                    raise ModelNotMatchingError(f"{s.name}: payload dtype {a.dtype} is not floating point")
                host[s.offset : s.offset + s.numel] = a.reshape(-1)
            else:
                extra[s.name] = torch.from_numpy(np.array(a))
        flat = torch.from_numpy(host)
        if self.device.type == "cuda":
            flat = flat.pin_memory().to(self.device, non_blocking=True)
        return FlatParams(flat, self.layout, extra)

    def check_parameters(self, params: Any) -> None:
        assert self.layout is not None
        if isinstance(params, FlatParams):
            if params.layout != self.layout:
                raise ModelNotMatchingError("flat model layout differs from the local model")
            return
        keys = list(params.keys())
        if keys != [s.name for s in self.layout.slots]:
            raise ModelNotMatchingError("state_dict keys differ from the local model")
        for s in self.layout.slots:
            if tuple(params[s.name].shape) != s.shape:
                raise ModelNotMatchingError(f"{s.name}: shape {tuple(params[s.name].shape)} != {s.shape}")

    def set_parameters(self, params: Any) -> None:
        assert self.flat is not None and self.layout is not None
        try:
            self.check_parameters(params)
        except ModelNotMatchingError:
            raise
        except Exception as e:
            raise ModelNotMatchingError(f"not matching models: {e}") from e
        with torch.no_grad():
            if isinstance(params, FlatParams):
                self.flat.copy_(params.flat.to(torch.float32), non_blocking=True)
                extra = params.extra
            else:
                views = self.layout.views(self.flat)
                extra = {}
                for s in self.layout.slots:
                    if s.in_arena:
                        views[s.name].copy_(torch.as_tensor(params[s.name]).to(torch.float32))
                    else:
                        extra[s.name] = torch.as_tensor(params[s.name])
            if extra:
                sd = self.model.state_dict()
                for k, v in extra.items():
                    sd[k].copy_(v.to(sd[k].device))
        self._params_written()

    # -------------------------------------------------------------- training
    def _make_optimizer(self) -> torch.optim.Optimizer:
        lr = float(getattr(self.model, "lr", 1e-3))
        if getattr(self.model, "optimizer", "adam") == "sgd":
            return torch.optim.SGD(self.model.parameters(), lr=lr)
        return torch.optim.Adam(self.model.parameters(), lr=lr)

    def _loss(self, logits: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        if logits.dim() == 3:  # language model: [b, t, vocab]
            return F.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), y.reshape(-1))
        return F.cross_entropy(logits.float(), y)

    def _autocast(self) -> Any:
        on = self.device.type == "cuda" and (Settings.AMP_BF16 or bool(getattr(self.model, "amp_bf16", False)))
        return torch.autocast(device_type=self.device.type, dtype=torch.bfloat16, enabled=on)

    def begin_round(self, round: Optional[int]) -> None:
        self._round_steps = 0

    def fit(self) -> None:
        if self.epochs <= 0 or self.model is None:
            return
        self._interrupt.clear()
        if self._optimizer is None or Settings.RESET_OPTIMIZER_EACH_ROUND:
            self._optimizer = self._make_optimizer()
        opt = self._optimizer
        self.model.train()
        step = 0
        for _ in range(self.epochs):
            for x, y in self.data.train_dataloader():
                if self._interrupt.is_set():
                    logger.info(self.addr, "fit interrupted")
                    return
                x, y = self._to_device(x, y)
                opt.zero_grad(set_to_none=True)
                with self._autocast():
                    loss = self._loss(self.model(x), y)
                loss.backward()
                opt.step()
                if step % self.LOG_EVERY == 0:
                    logger.log_metric(self.addr, "train_loss", float(loss.detach()), step=step)
                    self.d2h_bytes += 4
                step += 1
            # per-epoch validation, as Lightning does for the reference (mlp.py:82-91)
            val = self._score(self.data.val_dataloader()) if hasattr(self.data, "val_dataloader") else None
            if val is not None:
                logger.log_metric(self.addr, "val_loss", val[0], step=step)
                logger.log_metric(self.addr, "val_metric", val[1], step=step)
            self.model.train()
        self._round_steps = step

    def interrupt_fit(self) -> None:
        self._interrupt.set()

    @torch.no_grad()
    def _score(self, loader: Any) -> Optional[Tuple[float, float]]:
        """(mean loss, accuracy) over a loader; None when it is empty."""
        self.model.eval()
        loss_sum = torch.zeros((), device=self.device)
        correct = torch.zeros((), device=self.device)
        count = 0
        for x, y in loader:
            x, y = self._to_device(x, y)
            with self._autocast():
                logits = self.model(x)
            if logits.dim() == 3:
                logits, y = logits.reshape(-1, logits.shape[-1]), y.reshape(-1)
            loss_sum += F.cross_entropy(logits.float(), y, reduction="sum")
            correct += (logits.argmax(dim=1) == y).sum()
            count += y.numel()
        if count == 0:
            return None
        self.d2h_bytes += 8
        return float(loss_sum) / count, float(correct) / count

    def _to_device(self, x: torch.Tensor, y: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        if x.device != self.device:
            self.h2d_bytes += x.numel() * x.element_size() + y.numel() * y.element_size()
        return x.to(self.device, non_blocking=True), y.to(self.device, non_blocking=True)

    def evaluate(self) -> Dict[str, float]:
        if self.epochs <= 0 or self.model is None:
            return {}
        res = self._score(self.data.test_dataloader())
        if res is None:
            return {}
        out = {"test_loss": res[0], "test_metric": res[1]}
        for k, v in out.items():
            logger.log_metric(self.addr, k, v)
        return out

    def get_num_samples(self) -> Tuple[int, int]:
        if hasattr(self.data, "num_samples"):
            return self.data.num_samples()
        return len(self.data.train_dataloader().dataset), len(self.data.test_dataloader().dataset)
