"""Minimal stand-in for ``pytorch_lightning`` used ONLY to run the unmodified
reference (baseline/_ref) in an image that has no Lightning and no network.

Covers exactly what the reference touches (lightning_learner.py:180-236,
mlp.py / cnn.py): ``LightningModule`` (= nn.Module + ``self.log``),
``LightningDataModule``, and a ``Trainer`` with ``fit`` / ``test`` /
``should_stop`` doing a plain eager loop: forward -> loss -> backward ->
optimizer step, a validation pass per epoch, metrics handed to the logger every
``log_every_n_steps`` steps.  Lightning's sanity-check validation, progress bars,
checkpoints and callbacks are NOT emulated, which only makes the baseline
faster than real Lightning would be.
"""

from __future__ import annotations

import threading
from typing import Any, Dict, List, Optional

import torch
from torch import nn

__version__ = "0.0-shim"

# Bytes this shim moves between host and device, counted from the tensors it copies (batches and the
# module going to the device; metrics and the module coming back).  bench.py reads the counters to fill
# the reference arm's "e2e" block.  The reference's own copies (its encode/decode of model payloads)
# happen inside the reference and are not counted here.
COUNTERS: Dict[str, int] = {"h2d": 0, "d2h": 0}
_counters_lock = threading.Lock()  # the reference runs one trainer thread per node


def _count(key: str, n: int) -> None:
    with _counters_lock:
        COUNTERS[key] += n


def _nbytes(t: Any) -> int:
    return int(t.numel() * t.element_size()) if torch.is_tensor(t) else 0


def _module_bytes(m: nn.Module) -> int:
    return sum(_nbytes(p) for p in m.parameters()) + sum(_nbytes(b) for b in m.buffers())


def _to_dev(t: torch.Tensor, dev: torch.device, non_blocking: bool = False) -> torch.Tensor:
    if dev.type == "cuda" and not t.is_cuda:
        _count("h2d", _nbytes(t))
    return t.to(dev, non_blocking=non_blocking)


def _module_to(m: nn.Module, dev: Any) -> None:
    dev = torch.device(dev)
    try:
        cur = next(m.parameters()).device
    except StopIteration:
        cur = dev
    if cur.type != dev.type:
        _count("h2d" if dev.type == "cuda" else "d2h", _module_bytes(m))
    m.to(dev)


class LightningModule(nn.Module):
    def __init__(self, *a: Any, **kw: Any) -> None:
        super().__init__()
        self._shim_logged: Dict[str, List[Any]] = {}

    def log(self, name: str, value: Any, *a: Any, **kw: Any) -> None:
        self._shim_logged.setdefault(name, []).append(value.detach() if torch.is_tensor(value) else value)

    @property
    def device(self) -> torch.device:
        try:
            return next(self.parameters()).device
        except StopIteration:
            return torch.device("cpu")


class LightningDataModule:
    def __init__(self, *a: Any, **kw: Any) -> None:
        pass


def _pick_device(accelerator: str, data: Any) -> torch.device:
    if accelerator in ("cpu",) or not torch.cuda.is_available():
        return torch.device("cpu")
    hint = getattr(data, "device_hint", None)
    if hint is not None:
        return torch.device(hint)
    return torch.device(f"cuda:{torch.cuda.current_device()}")


def _mean(vals: List[Any]) -> float:
    if not vals:
        return float("nan")
    _count("d2h", sum(_nbytes(v) for v in vals if torch.is_tensor(v) and v.is_cuda))
    return float(torch.stack([torch.as_tensor(v, dtype=torch.float32).cpu() for v in vals]).mean())


class Trainer:
    def __init__(
        self,
        max_epochs: int = 1,
        accelerator: str = "auto",
        logger: Any = None,
        enable_checkpointing: bool = False,
        enable_model_summary: bool = False,
        log_every_n_steps: int = 50,
        **kw: Any,
    ) -> None:
        self.max_epochs = max_epochs
        self.accelerator = accelerator
        self.logger = logger if logger not in (False, None) else None
        self.log_every_n_steps = log_every_n_steps
        self.should_stop = False
        self.global_step = 0

    def fit(self, model: LightningModule, datamodule: Any) -> None:
        dev = _pick_device(self.accelerator, datamodule)
        _module_to(model, dev)
        opt = model.configure_optimizers()
        for epoch in range(self.max_epochs):
            model.train()
            for i, (x, y) in enumerate(datamodule.train_dataloader()):
                if self.should_stop:
                    break
                x, y = _to_dev(x, dev, True), _to_dev(y, dev, True)
                model._shim_logged = {}
                loss = model.training_step((x, y), i)
                opt.zero_grad()
                loss.backward()
                opt.step()
                self.global_step += 1
                if self.logger is not None and self.log_every_n_steps and self.global_step % self.log_every_n_steps == 0:
                    m = {k: _mean(v) for k, v in model._shim_logged.items()}
                    m["epoch"] = epoch
                    self.logger.log_metrics(m, self.global_step)
            if self.should_stop:
                break
            model.eval()
            model._shim_logged = {}
            with torch.no_grad():
                for i, (x, y) in enumerate(datamodule.val_dataloader()):
                    model.validation_step((_to_dev(x, dev), _to_dev(y, dev)), i)
            if self.logger is not None and model._shim_logged:
                m = {k: _mean(v) for k, v in model._shim_logged.items()}
                m["epoch"] = epoch
                self.logger.log_metrics(m, self.global_step)
        _module_to(model, "cpu")  # Lightning's teardown moves the module back to the host

    def test(self, model: LightningModule, datamodule: Any, verbose: bool = False) -> List[Dict[str, float]]:
        dev = _pick_device(self.accelerator, datamodule)
        _module_to(model, dev)
        model.eval()
        model._shim_logged = {}
        with torch.no_grad():
            for i, (x, y) in enumerate(datamodule.test_dataloader()):
                model.test_step((_to_dev(x, dev), _to_dev(y, dev)), i)
        out = {k: _mean(v) for k, v in model._shim_logged.items()}
        _module_to(model, "cpu")
        return [out]
