"""Transport-neutral envelopes.

Field-for-field the protobuf messages of the reference
(p2pfl/communication/grpc/proto/node.proto:26-48) so the gRPC transport can
marshal them byte-compatibly, while the in-memory and NVLink transports pass
the objects (or device handles) by reference.
"""

from __future__ import annotations

import random
from dataclasses import dataclass, field, replace
from typing import Any, List


def new_message_id() -> int:
    """63-bit id used only for duplicate suppression (reference hashes
    cmd+args+time+randint with Python ``hash``; grpc_client.py:66-72)."""
    return random.getrandbits(63)


@dataclass
class Message:
    """Small control message, flooded with a TTL."""

    source: str
    cmd: str
    args: List[str] = field(default_factory=list)
    round: int = -1  # -1 = "no round"
    ttl: int = 1
    hash: int = 0

    def relay_copy(self) -> "Message":
        """Copy-on-relay: TTL is decremented on the copy, never on the object
        other receivers still hold (fixes SURVEY Q6)."""
        return replace(self, ttl=self.ttl - 1, args=list(self.args))


@dataclass
class Weights:
    """Model-carrying message (never relayed, never deduplicated).

    ``weights`` is ``bytes`` on the host path (pickled list of ndarrays) or a
    device payload object on the device path (see
    ``p2pfl_b200.parallel.payload``)."""

    source: str
    cmd: str
    round: int
    weights: Any
    contributors: List[str] = field(default_factory=list)
    weight: int = 1
