"""Device payloads: what a ``Weights`` message carries on the device data plane.

Host path:   ``Weights.weights`` = pickled ndarrays (bytes), byte-compatible with the reference.
Device path: ``Weights.weights`` = :class:`DevicePayload` while inside the sending process; on the
             control wire it is a ~200-byte *reference* (``DeviceRef``) and the receiver pulls the
             tensor over NVLink from the sender's arena (or copies device-to-device in-process).
"""

from __future__ import annotations

import json
from dataclasses import asdict, dataclass
from typing import Any, Dict

from p2pfl_b200.learning.params import FlatParams

MAGIC = b"P2PFL-DEVREF1:"


class DevicePayload:
    """A model that stays on the GPU.  ``owner`` is the sending learner (gives access to
    its arena for staging when the receiver lives in another process)."""

    def __init__(self, params: FlatParams, owner: Any) -> None:
        self.params = params
        self.owner = owner


@dataclass
class DeviceRef:
    """Control-wire form of a device payload."""

    addr: str  # sender (owner of the arena to pull from)
    region: str  # "params" | "outbox"
    layout_hash: str
    n: int
    extra: Dict[str, Any]  # out-of-band integer buffers (tiny), as lists
    nonce: int

    def to_bytes(self) -> bytes:
        return MAGIC + json.dumps(asdict(self)).encode()

    @staticmethod
    def is_ref(data: Any) -> bool:
        return isinstance(data, (bytes, bytearray)) and bytes(data[: len(MAGIC)]) == MAGIC

    @classmethod
    def from_bytes(cls, data: bytes) -> "DeviceRef":
        return cls(**json.loads(bytes(data)[len(MAGIC) :].decode()))
