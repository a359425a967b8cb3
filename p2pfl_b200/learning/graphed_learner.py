"""Graphed eager learner: PyTorch numerics (fp32 autograd, ``torch.optim.Adam``), no host work per step.

What the reference's learner does per batch - DataLoader collation, one ``.to(device)`` per tensor,
~30 kernel launches of forward / loss / backward / optimizer issued from Python
(p2pfl/learning/pytorch/lightning_learner.py:180-207 through ``Trainer.fit``) - costs about a
millisecond of host time for an MLP step whose device time is tens of microseconds.  This learner keeps
the same operators and the same precision (strict fp32: it is the precision-parity arm of ``bench.py``)
and removes the host from the loop the way the fused learner does:

* the splits are uploaded once per round from pinned host memory (or stay resident), shuffled by one
  ``randperm`` + gather on the device;
* ONE train step (forward, cross-entropy, backward, capturable Adam) is captured in a CUDA graph on
  static input buffers and replayed per batch (two small copies + one graph launch per step);
* validation and evaluation run over the resident split in a few large batches.

It is a library path (cuBLAS / ATen kernels) - the product path is ``B200Learner`` - and exists for two
reasons: models without a fused plan get most of the launch overhead back, and the benchmark has an arm
whose dtype equals the reference's.  Without CUDA it behaves exactly like ``TorchLearner``.
"""

from __future__ import annotations

from typing import Any, Dict, Optional, Tuple

import torch
import torch.nn.functional as F

from p2pfl_b200.learning.torch_learner import TorchLearner
from p2pfl_b200.management.logger import logger
from p2pfl_b200.settings import Settings


class GraphedTorchLearner(TorchLearner):
    EVAL_CHUNK = 8192

    def __init__(self, model: Any, data: Any, self_addr: str = "unknown-node", epochs: int = 1, device: Optional[str] = None,
                 data_mode: str = "stream", shuffle: bool = True) -> None:  # fmt: skip
        self.data_mode = data_mode  # "stream": H2D of the splits every round; "resident": once
        self.shuffle = shuffle
        self._graph: Optional[Dict[str, Any]] = None
        self._dev_split: Dict[str, Tuple[torch.Tensor, torch.Tensor]] = {}
        self._stream: Optional[torch.cuda.Stream] = None
        super().__init__(model, data, self_addr, epochs, device)

    # ------------------------------------------------------------------ plumbing
    @property
    def graphed(self) -> bool:
        return self.device.type == "cuda" and hasattr(self.data, "tensors")

    @property
    def stream(self) -> Any:
        """The stream this learner computes on (``bench.py`` times on it)."""
        if self.device.type != "cuda":
            return None
        if self._stream is None:
            self._stream = torch.cuda.Stream(device=self.device)
        return self._stream

    def set_model(self, model: Any) -> None:
        super().set_model(model)
        self._graph = None  # parameter storage moved: the captured step is stale

    def set_data(self, data: Any) -> None:
        super().set_data(data)
        self._dev_split = {}
        self._graph = None

    def _split(self, name: str, refresh: bool) -> Tuple[torch.Tensor, torch.Tensor]:
        """Device copy of a split; re-uploaded (from pinned memory, counted) when ``refresh``."""
        have = self._dev_split.get(name)
        if have is not None and not refresh:
            return have
        src = self.data.pinned_tensors(name) if hasattr(self.data, "pinned_tensors") else self.data.tensors(name)
        x, y = src
        if have is None:
            have = (torch.empty(x.shape, dtype=x.dtype, device=self.device), torch.empty(y.shape, dtype=torch.long, device=self.device))
            self._dev_split[name] = have
        have[0].copy_(x, non_blocking=True)
        have[1].copy_(y.to(torch.long) if y.dtype != torch.long else y, non_blocking=True)
        self.h2d_bytes += x.numel() * x.element_size() + y.numel() * 8
        return have

    # ------------------------------------------------------------------ the captured step
    def _make_optimizer(self) -> torch.optim.Optimizer:
        if not self.graphed:
            return super()._make_optimizer()
        lr = float(getattr(self.model, "lr", 1e-3))
        if getattr(self.model, "optimizer", "adam") == "sgd":
            return torch.optim.SGD(self.model.parameters(), lr=lr)
        return torch.optim.Adam(self.model.parameters(), lr=lr, capturable=True, foreach=True)

    def _reset_optimizer_state(self) -> None:
        assert self._optimizer is not None
        for st in self._optimizer.state.values():
            for v in st.values():
                if torch.is_tensor(v):
                    v.zero_()

    def _build_step(self, x: torch.Tensor, y: torch.Tensor, B: int) -> Dict[str, Any]:
        """Capture forward + loss + backward + optimizer step for batches of ``B`` rows."""
        assert self.flat is not None
        self._optimizer = self._make_optimizer()
        opt = self._optimizer
        xs = torch.zeros((B, *x.shape[1:]), dtype=x.dtype, device=self.device)
        ys = torch.zeros((B,), dtype=torch.long, device=self.device)
        xs.copy_(x[:B])
        ys.copy_(y[:B])
        loss_out = torch.zeros((), device=self.device)
        keep = self.flat.detach().clone()
        self.model.train()

        def step() -> torch.Tensor:
            opt.zero_grad(set_to_none=True)
            with self._autocast():
                loss = self._loss(self.model(xs), ys)
            loss.backward()
            opt.step()
            return loss

        for _ in range(3):  # warm-up outside the capture: lazy initialisation of optimizer state, cuBLAS handles, autograd
            step()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=self.stream, capture_error_mode="thread_local"):
            loss_out.copy_(step().detach())
        # the warm-up steps trained on the first batch: put the model and the optimizer back
        self.flat.copy_(keep)
        self._reset_optimizer_state()
        return {"graph": g, "x": xs, "y": ys, "loss": loss_out, "B": B, "step": step}

    # ------------------------------------------------------------------ fit / evaluate
    def fit(self) -> None:
        if not self.graphed:
            return super().fit()
        if self.epochs <= 0 or self.model is None:
            return
        self._interrupt.clear()
        B = int(getattr(self.data, "batch_size", 32))
        self.stream.wait_stream(torch.cuda.current_stream(self.device))  # whoever wrote the parameters last
        with torch.cuda.stream(self.stream):
            x, y = self._split("train", refresh=self.data_mode == "stream")
            n = x.shape[0]
            if self._graph is None or self._graph["B"] != min(B, n):
                self._graph = self._build_step(x, y, min(B, n))
            elif Settings.RESET_OPTIMIZER_EACH_ROUND:
                self._reset_optimizer_state()  # the reference builds a fresh Trainer (=> fresh Adam) per round (Q16)
            gr = self._graph
            B = gr["B"]
            nb_full = n // B
            n_log = (nb_full + self.LOG_EVERY - 1) // self.LOG_EVERY
            log = torch.zeros(max(1, n_log), device=self.device)
            step = 0
            for e in range(self.epochs):
                if self._interrupt.is_set():
                    logger.info(self.addr, "fit interrupted")
                    break
                if self.shuffle:
                    perm = torch.randperm(n, device=self.device)
                    xe, ye = x.index_select(0, perm), y.index_select(0, perm)
                else:
                    xe, ye = x, y
                self.model.train()
                for i in range(nb_full):
                    gr["x"].copy_(xe[i * B : (i + 1) * B])
                    gr["y"].copy_(ye[i * B : (i + 1) * B])
                    gr["graph"].replay()
                    if e == self.epochs - 1 and i % self.LOG_EVERY == 0:
                        log[i // self.LOG_EVERY].copy_(gr["loss"])
                    step += 1
                if n - nb_full * B > 0:  # the short last batch runs eagerly (one step per epoch)
                    self._eager_step(xe[nb_full * B :], ye[nb_full * B :])
                    step += 1
                val = self._score_resident("val")
                if val is not None:
                    logger.log_metric(self.addr, "val_loss", val[0], step=step)
                    logger.log_metric(self.addr, "val_metric", val[1], step=step)
            losses = log.cpu()
            self.d2h_bytes += losses.numel() * 4
        for j in range(n_log):
            logger.log_metric(self.addr, "train_loss", float(losses[j]), step=j * self.LOG_EVERY)
        self._round_steps = step

    def _eager_step(self, xb: torch.Tensor, yb: torch.Tensor) -> None:
        assert self._optimizer is not None
        self._optimizer.zero_grad(set_to_none=True)
        with self._autocast():
            loss = self._loss(self.model(xb), yb)
        loss.backward()
        self._optimizer.step()

    @torch.no_grad()
    def _score_resident(self, split: str) -> Optional[Tuple[float, float]]:
        if not hasattr(self.data, "tensors"):
            return None
        try:
            x, y = self._split(split, refresh=self.data_mode == "stream")
        except KeyError:
            return None
        if x.shape[0] == 0:
            return None
        self.model.eval()
        acc = torch.zeros(2, device=self.device)
        for i in range(0, x.shape[0], self.EVAL_CHUNK):
            xb, yb = x[i : i + self.EVAL_CHUNK], y[i : i + self.EVAL_CHUNK]
            with self._autocast():
                logits = self.model(xb)
            if logits.dim() == 3:
                logits, yb = logits.reshape(-1, logits.shape[-1]), yb.reshape(-1)
            acc[0] += F.cross_entropy(logits.float(), yb, reduction="sum")
            acc[1] += (logits.argmax(dim=1) == yb).sum()
        out = acc.cpu()
        self.d2h_bytes += 8
        count = y.numel()
        return float(out[0]) / count, float(out[1]) / count

    def evaluate(self) -> Dict[str, float]:
        if not self.graphed:
            return super().evaluate()
        if self.epochs <= 0 or self.model is None:
            return {}
        self.stream.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.stream):
            res = self._score_resident("test")
        if res is None:
            return {}
        out = {"test_loss": res[0], "test_metric": res[1]}
        for k, v in out.items():
            logger.log_metric(self.addr, k, v)
        return out

    # parameters written from outside (aggregation) land on other streams: order them before ours
    def set_parameters(self, params: Any) -> None:
        if not self.graphed:
            return super().set_parameters(params)
        with torch.cuda.stream(self.stream):
            super().set_parameters(params)
        self.stream.synchronize()
