"""Baseline exchange: the round's models through a library collective.

"NCCL all-gather + average" is what one would write without the fused kernel: every trainer
contributes ``n_i * model_i`` (and ``n_i``), everybody gathers everything, sums in rank order and divides
by the gathered total.  ``bench/gossip_bw.py`` times that collective alone next to the fused kernel;
this module makes it a FULL-ROUND arm: the same node, stages, learner and control plane, only
``plan_fused_exchange`` hands the stage a plan whose ``launch`` enqueues the collective on the
learner's stream instead of ``p2pfl_gossip_fused`` (``bench.py --exchange nccl``).

It covers one case on purpose - the benchmark's: every rank of the process group hosts exactly one
node and all of them train.  Anything else returns no plan and the round is exchanged through
``add_model`` messages.  The result is bit-identical on every rank (same inputs, same summation order).
On CPU tensors it runs over ``gloo`` (tests/test_collective_exchange.py).
"""

from __future__ import annotations

import contextlib
from typing import Any, List, Optional, Tuple, Type

import torch
import torch.distributed as dist


class AllGatherAverage:
    """Plan object with the interface TrainStage expects of a fused exchange
    (``can_prelaunch`` / ``launch`` / ``finish`` / ``run``)."""

    can_prelaunch = True

    def __init__(self, state: Any, learner: Any, train_set: List[str], group: Any = None) -> None:
        self.state, self.learner, self.train_set, self.group = state, learner, list(train_set), group
        self.complete = False
        self._total: Optional[torch.Tensor] = None
        flat = learner.flat
        world = dist.get_world_size(group)
        self._send = torch.empty(flat.numel() + 1, dtype=torch.float32, device=flat.device)
        self._recv = [torch.empty_like(self._send) for _ in range(world)]

    def _stream(self) -> Any:
        s = getattr(self.learner, "stream", None)
        return torch.cuda.stream(s) if s is not None and self.learner.flat.is_cuda else contextlib.nullcontext()

    def launch(self, seq: int) -> None:
        learner = self.learner
        flat = learner.flat
        w = float(learner.get_num_samples()[0])
        with self._stream(), torch.no_grad():
            self._send[:-1].copy_(flat)
            self._send[:-1].mul_(w)
            self._send[-1] = w
            dist.all_gather(self._recv, self._send, group=self.group)  # ordered after the training on this stream
            acc = self._recv[0].clone()
            for t in self._recv[1:]:  # fixed (rank) order: identical bits everywhere
                acc.add_(t)
            flat.copy_(acc[:-1] / acc[-1])
            self._total = acc[-1:].clone()
            written = getattr(learner, "_params_written", None)
            if written is not None:
                written()  # refresh the bf16 working copies of the fused learner

    def finish(self) -> Optional[Tuple[List[str], int]]:
        assert self._total is not None, "finish() before launch()"
        total = int(round(float(self._total.cpu())))  # the D2H read orders the host behind the collective
        self.complete = True
        return list(self.train_set), total

    def run(self, seq: int) -> Optional[Tuple[List[str], int]]:
        self.launch(seq)
        return self.finish()


def with_collective_exchange(base: Type) -> Type:
    """``base`` (a protocol class) with the round's exchange done by all-gather + average."""

    class CollectiveExchangeProtocol(base):  # type: ignore[misc,valid-type]
        exchange = "all-gather + average (library collective)"

        def plan_fused_exchange(self, state: Any, learner: Any, wait: Optional[float] = None) -> Any:
            if not dist.is_available() or not dist.is_initialized() or getattr(learner, "flat", None) is None:
                return None
            train_set = sorted(state.train_set)
            # one node per rank, all of them training: the only case this baseline covers
            if state.addr not in train_set or len(train_set) != dist.get_world_size() or len(train_set) < 2:
                return None
            stats = getattr(self, "stats", None)
            if stats is not None:
                stats["collective"] = stats.get("collective", 0) + 1
            return AllGatherAverage(state, learner, train_set)

    CollectiveExchangeProtocol.__name__ = f"Collective{base.__name__}"
    return CollectiveExchangeProtocol
