"""Federated averaging: ``sum_i w_i * m_i / sum_i w_i``
(reference: p2pfl/learning/aggregators/fedavg.py:31-60).

Three paths, chosen by the type of the models:

* ``FlatParams`` on CUDA  -> one fused sm_100a kernel over the flat vectors
  (``p2pfl_b200.ops.fedavg_flat``), fp32 accumulate;
* ``FlatParams`` on CPU   -> vectorised torch ops on the flat vectors;
* ``{name: tensor}``      -> the reference's per-tensor loop; this is the CPU
  oracle the kernels are tested against.

Integer entries are not averaged (true division would turn them into floats,
SURVEY Q14): the value of the heaviest contributor is kept.
"""

from __future__ import annotations

from collections import OrderedDict
from typing import Any, Dict, Tuple

import torch

from p2pfl_b200.exceptions import ModelNotMatchingError, NoModelsToAggregateError
from p2pfl_b200.learning.aggregators.aggregator import Aggregator
from p2pfl_b200.learning.params import FlatParams


def fedavg_state_dicts(models: Dict[str, Tuple[Dict[str, torch.Tensor], int]]) -> "OrderedDict[str, torch.Tensor]":
    """Reference arithmetic on state-dicts (the oracle)."""
    entries = list(models.values())
    total = float(sum(w for _, w in entries))
    if total == 0:
        raise NoModelsToAggregateError("total weight is zero")
    first = entries[0][0]
    heaviest = max(entries, key=lambda e: e[1])[0]
    out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for name, ref in first.items():
        if not torch.is_tensor(ref):
            ref = torch.as_tensor(ref)
        if ref.is_floating_point():
            acc = torch.zeros_like(ref, dtype=torch.float32)
            for m, w in entries:
                acc += torch.as_tensor(m[name]).to(torch.float32) * float(w)
            out[name] = (acc / total).to(ref.dtype)
        else:
            out[name] = torch.as_tensor(heaviest[name]).clone()
    return out


def fedavg_flat(models: Dict[str, Tuple[FlatParams, int]]) -> FlatParams:
    entries = list(models.values())
    first = entries[0][0]
    for m, _ in entries[1:]:
        if m.layout != first.layout:
            raise ModelNotMatchingError("cannot average models with different layouts")
    total = float(sum(w for _, w in entries))
    if total == 0:
        raise NoModelsToAggregateError("total weight is zero")
    heaviest = max(entries, key=lambda e: e[1])[0]
    flats = [m.flat for m, _ in entries]
    weights = [float(w) / total for _, w in entries]
    if first.flat.is_cuda:
        from p2pfl_b200 import ops

        out = ops.fedavg_flat(flats, weights, out_dtype=torch.float32)
    else:
        out = torch.zeros(first.layout.total, dtype=torch.float32)
        for f, w in zip(flats, weights):
            out.add_(f.to(torch.float32), alpha=w)
    return FlatParams(out, first.layout, {k: v.clone() for k, v in heaviest.extra.items()})


class FedAvg(Aggregator):
    """``FedAvg().aggregate({key: (model, weight)})`` is usable stand-alone
    (reference: test/learning_test.py:54-59)."""

    def __init__(self, node_name: str = "unknown") -> None:
        super().__init__(node_name=node_name)

    def aggregate(self, models: Dict[str, Tuple[Any, int]]) -> Any:
        if len(models) == 0:
            raise NoModelsToAggregateError(f"({self.node_name}) trying to aggregate models when there is no models")
        if all(isinstance(m, FlatParams) for m, _ in models.values()):
            return fedavg_flat(models)  # type: ignore[arg-type]
        as_dicts = {
            k: ((m.state_dict() if isinstance(m, FlatParams) else m), w) for k, (m, w) in models.items()
        }
        return fedavg_state_dicts(as_dicts)
