"""Flat parameter representation.

The device wire format of this framework is *the model as one contiguous
vector*: every floating-point ``state_dict`` entry laid out back to back, each
starting on a 128-byte boundary, in ``state_dict`` order (SURVEY §5.8 "flat
parameter arena").  Integer buffers (BatchNorm ``num_batches_tracked``) are
kept out of band and are not averaged.

``ParamLayout`` is the static description, ``FlatParams`` a vector plus its
layout.  A layout hash is compared before a peer's vector is accepted, which
replaces the reference's "zip with local keys and hope" check (Q20).
"""

from __future__ import annotations

import hashlib
from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, List, Mapping, Optional, Tuple

import torch

ALIGN_BYTES = 128
_ALIGN_ELEMS = ALIGN_BYTES // 4  # offsets are in fp32 elements


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


@dataclass(frozen=True)
class ParamSlot:
    name: str
    shape: Tuple[int, ...]
    dtype: torch.dtype  # dtype of the state_dict entry (wire dtype on the host path)
    offset: int  # element offset in the flat vector, -1 for out-of-band entries
    numel: int

    @property
    def in_arena(self) -> bool:
        return self.offset >= 0


class ParamLayout:
    """Ordered slots + total (padded) length of the flat vector."""

    def __init__(self, slots: List[ParamSlot], total: int) -> None:
        self.slots = slots
        self.total = total
        self.by_name: Dict[str, ParamSlot] = {s.name: s for s in slots}
        h = hashlib.sha1()
        for s in slots:
            h.update(f"{s.shape}|{s.dtype}|{s.offset};".encode())
        self.hash = h.hexdigest()[:16]

    @classmethod
    def from_state_dict(cls, sd: Mapping[str, torch.Tensor]) -> "ParamLayout":
        slots: List[ParamSlot] = []
        off = 0
        for name, t in sd.items():
            if t.is_floating_point():
                slots.append(ParamSlot(name, tuple(t.shape), t.dtype, off, t.numel()))
                off = _round_up(off + t.numel(), _ALIGN_ELEMS)
            else:
                slots.append(ParamSlot(name, tuple(t.shape), t.dtype, -1, t.numel()))
        return cls(slots, max(off, _ALIGN_ELEMS))

    @property
    def num_params(self) -> int:
        return sum(s.numel for s in self.slots if s.in_arena)

    def views(self, flat: torch.Tensor) -> "OrderedDict[str, torch.Tensor]":
        """Per-tensor views into ``flat`` (arena entries only)."""
        out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        for s in self.slots:
            if s.in_arena:
                out[s.name] = flat[s.offset : s.offset + s.numel].view(s.shape)
        return out

    def __eq__(self, other: object) -> bool:
        return isinstance(other, ParamLayout) and other.hash == self.hash and other.total == self.total

    def __hash__(self) -> int:
        return hash((self.hash, self.total))


class FlatParams:
    """A model as one vector.  ``extra`` carries the out-of-band integer buffers."""

    __slots__ = ("flat", "layout", "extra")

    def __init__(self, flat: torch.Tensor, layout: ParamLayout, extra: Optional[Dict[str, torch.Tensor]] = None):
        if flat.dim() != 1 or flat.numel() != layout.total:
            raise ValueError(f"flat vector has {tuple(flat.shape)}, layout wants ({layout.total},)")
        self.flat = flat
        self.layout = layout
        self.extra = extra or {}

    def state_dict(self) -> "OrderedDict[str, torch.Tensor]":
        views = self.layout.views(self.flat)
        out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        for s in self.layout.slots:
            out[s.name] = views[s.name] if s.in_arena else self.extra[s.name]
        return out

    def clone(self) -> "FlatParams":
        return FlatParams(self.flat.clone(), self.layout, {k: v.clone() for k, v in self.extra.items()})
