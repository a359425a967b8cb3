"""Device arenas: the memory the gossip data plane works on.

One arena per federated node (SURVEY §5.8 "flat parameter arena"), allocated
by the C++ runtime with ``cudaMalloc`` so it can be exported over CUDA IPC and
mapped into peer GPUs / peer processes.  Layout (every region 256-byte aligned):

    params   fp32[n]                       the live model (the learner's tensors are views)
    outbox   fp32[n]                       staging for models peers pull (partial aggregates)
    inbox    wire[2 banks][slots][n]       one slot per possible contributor
    flags    u32 [2 banks][slots][tiles]   arrival flag per (slot, tile), holds the exchange seq
    status   i32 [4]                       written by the reduce kernels
    abort    u32 [1]                       host watchdog -> spinning kernels

Two inbox banks (``seq & 1``) let a peer that is one round ahead push its next
model while this node still reads the previous one.  The geometry is a pure
function of ``(n, slots, tile, wire dtype)``, so a peer needs only the base
pointer (IPC handle) to address any region.

No counterpart in the reference (models live in host ``state_dict``s there).
"""

from __future__ import annotations

import threading
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from p2pfl_b200 import ops


def _align(x: int, a: int = 256) -> int:
    return (x + a - 1) // a * a


class _RawCudaBuffer:
    """Exposes a raw device pointer to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, nbytes: int, typestr: str, shape: Tuple[int, ...]) -> None:
        self.__cuda_array_interface__ = {
            "shape": shape,
            "typestr": typestr,
            "data": (ptr, False),
            "version": 3,
            "strides": None,
        }
        self._nbytes = nbytes


def tensor_from_ptr(ptr: int, numel: int, dtype: torch.dtype, device: int) -> torch.Tensor:
    """Zero-copy torch view of runtime-owned device memory (bf16 goes through int16)."""
    if dtype == torch.bfloat16:
        return tensor_from_ptr(ptr, numel, torch.int16, device).view(torch.bfloat16)
    typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.int16: "<i2", torch.uint8: "|u1"}[dtype]
    itemsize = torch.empty(0, dtype=dtype).element_size()
    with torch.cuda.device(device):
        return torch.as_tensor(_RawCudaBuffer(ptr, numel * itemsize, typestr, (numel,)), device=f"cuda:{device}")


@dataclass(frozen=True)
class ArenaGeometry:
    n: int
    slots: int
    tile: int
    wire_bf16: bool

    @property
    def wire_bytes(self) -> int:
        return 2 if self.wire_bf16 else 4

    @property
    def num_tiles(self) -> int:
        return (self.n + self.tile - 1) // self.tile

    @property
    def off_params(self) -> int:
        return 0

    @property
    def off_outbox(self) -> int:
        return _align(self.n * 4)

    @property
    def off_inbox(self) -> int:
        return 2 * _align(self.n * 4)

    @property
    def slot_bytes(self) -> int:
        return _align(self.n * self.wire_bytes)

    @property
    def off_flags(self) -> int:
        return self.off_inbox + 2 * self.slots * self.slot_bytes

    @property
    def flag_row_bytes(self) -> int:
        return _align(self.num_tiles * 4)

    @property
    def off_status(self) -> int:
        return self.off_flags + 2 * self.slots * self.flag_row_bytes

    @property
    def off_abort(self) -> int:
        return self.off_status + 256

    @property
    def total_bytes(self) -> int:
        return self.off_abort + 256

    def slot_offset(self, slot: int, bank: int) -> int:
        return self.off_inbox + (bank * self.slots + slot) * self.slot_bytes

    def flag_offset(self, slot: int, bank: int) -> int:
        return self.off_flags + (bank * self.slots + slot) * self.flag_row_bytes

    def key(self) -> str:
        return f"{self.n}:{self.slots}:{self.tile}:{int(self.wire_bf16)}"

    @classmethod
    def from_key(cls, key: str) -> "ArenaGeometry":
        n, s, t, w = key.split(":")
        return cls(int(n), int(s), int(t), bool(int(w)))


class DeviceArena:
    """Owning handle of one arena."""

    def __init__(self, n: int, slots: int, tile: int = 16384, wire_bf16: bool = True, device: int = 0) -> None:
        if n % 8:
            raise ValueError("arena length must be a multiple of 8 elements")
        self.geo = ArenaGeometry(n, slots, tile, wire_bf16)
        self.device = device
        self.base = ops.rt_alloc(device, self.geo.total_bytes)
        self.params = tensor_from_ptr(self.base + self.geo.off_params, n, torch.float32, device)
        self.outbox = tensor_from_ptr(self.base + self.geo.off_outbox, n, torch.float32, device)
        self.status = tensor_from_ptr(self.base + self.geo.off_status, 4, torch.int32, device)
        self.abort = tensor_from_ptr(self.base + self.geo.off_abort, 1, torch.int32, device)
        self._closed = False

    # geometry shortcuts
    @property
    def n(self) -> int:
        return self.geo.n

    @property
    def num_tiles(self) -> int:
        return self.geo.num_tiles

    def slot_ptr(self, slot: int, bank: int) -> int:
        return self.base + self.geo.slot_offset(slot, bank)

    def flag_ptr(self, slot: int, bank: int) -> int:
        return self.base + self.geo.flag_offset(slot, bank)

    def slot_tensor(self, slot: int, bank: int) -> torch.Tensor:
        dt = torch.bfloat16 if self.geo.wire_bf16 else torch.float32
        return tensor_from_ptr(self.slot_ptr(slot, bank), self.n, dt, self.device)

    def flags_tensor(self, slot: int, bank: int) -> torch.Tensor:
        return tensor_from_ptr(self.flag_ptr(slot, bank), self.num_tiles, torch.int32, self.device)

    def ipc_handle(self) -> bytes:
        return ops.rt_ipc_export(self.device, self.base)

    def clear_arrivals(self) -> None:
        """Zero both banks of arrival flags, the status words and the abort word.

        Readiness of a tile is ``flag - seq >= 0`` (signed), so a flag left over from an earlier
        experiment must never survive into one whose sequence numbers start lower (a reference
        initiator always starts at 0; the wall-clock base wraps).  Called when the arena goes back to
        the pool: by then the last round's reduce has consumed every push addressed to this arena and
        nobody pushes again before the next experiment announces itself."""
        if self._closed:
            return
        geo = self.geo
        nbytes = geo.off_abort + 256 - geo.off_flags
        # stream-ordered and waited for on that stream only: a device-wide synchronize is refused by
        # CUDA while any other node of the process captures a graph
        with torch.cuda.device(self.device):
            s = torch.cuda.current_stream(self.device)
            tensor_from_ptr(self.base + geo.off_flags, nbytes, torch.uint8, self.device).zero_()
            s.synchronize()

    def close(self) -> None:
        if not self._closed:
            self._closed = True
            # the owner has synchronised the streams that used the arena; here only the caller's stream
            # (never the whole device: see clear_arrivals)
            torch.cuda.current_stream(self.device).synchronize()
            self.params = self.outbox = self.status = self.abort = None  # type: ignore[assignment]
            ops.rt_free(self.device, self.base)

    def __del__(self) -> None:  # best effort
        try:
            self.close()
        except Exception:
            pass


class ArenaPool:
    """Process-wide pool so that a node's arena (and with it the IPC handle its peers
    hold) survives from one experiment to the next instead of being re-allocated."""

    _lock = threading.Lock()
    _free: Dict[Tuple[str, int, str], DeviceArena] = {}

    @classmethod
    def acquire(cls, owner: str, n: int, slots: int, tile: int, wire_bf16: bool, device: int) -> DeviceArena:
        key = (owner, device, ArenaGeometry(n, slots, tile, wire_bf16).key())
        with cls._lock:
            arena = cls._free.pop(key, None)
        if arena is None:
            arena = DeviceArena(n, slots, tile, wire_bf16, device)
        arena.pool_key = key  # type: ignore[attr-defined]
        return arena

    @classmethod
    def release(cls, arena: DeviceArena) -> None:
        key = getattr(arena, "pool_key", None)
        if key is None or arena._closed:
            arena.close()
            return
        arena.clear_arrivals()
        with cls._lock:
            old = cls._free.pop(key, None)
            cls._free[key] = arena
        if old is not None and old is not arena:
            old.close()

    @classmethod
    def drain(cls) -> None:
        with cls._lock:
            arenas, cls._free = list(cls._free.values()), {}
        for a in arenas:
            a.close()


class PeerArena:
    """A peer's arena as seen from this process: a base pointer + its geometry."""

    def __init__(self, base: int, geo: ArenaGeometry, device: int, ipc_opened_on: Optional[int] = None) -> None:
        self.base, self.geo, self.device = base, geo, device
        self._ipc_dev = ipc_opened_on

    def slot_ptr(self, slot: int, bank: int) -> int:
        return self.base + self.geo.slot_offset(slot, bank)

    def flag_ptr(self, slot: int, bank: int) -> int:
        return self.base + self.geo.flag_offset(slot, bank)

    def params_ptr(self) -> int:
        return self.base + self.geo.off_params

    def region_ptr(self, region: str) -> int:
        return self.base + (self.geo.off_outbox if region == "outbox" else self.geo.off_params)

    def close(self) -> None:
        if self._ipc_dev is not None:
            dev, self._ipc_dev = self._ipc_dev, None
            _IpcCache.release(dev, self.base)

    @classmethod
    def open_ipc(cls, handle: bytes, geo: ArenaGeometry, local_device: int, peer_device: int) -> "PeerArena":
        return cls(_IpcCache.open(local_device, handle), geo, peer_device, ipc_opened_on=local_device)


class _IpcCache:
    """One ``cudaIpcOpenMemHandle`` per (peer allocation, local device) and PROCESS.

    Several federated nodes of one process (BASELINE config 4: eight per GPU) all talk to the same
    remote arenas; CUDA maps an exported allocation once per process - a second open of the same
    handle fails - so the mapping is shared and reference counted."""

    _lock = threading.Lock()
    _by_handle: Dict[Tuple[int, bytes], List[int]] = {}  # (device, handle) -> [base, refs]
    _by_base: Dict[Tuple[int, int], bytes] = {}

    @classmethod
    def open(cls, device: int, handle: bytes) -> int:
        with cls._lock:
            ent = cls._by_handle.get((device, handle))
            if ent is None:
                base = ops.rt_ipc_open(device, handle)
                ent = cls._by_handle[(device, handle)] = [base, 0]
                cls._by_base[(device, base)] = handle
            ent[1] += 1
            return ent[0]

    @classmethod
    def release(cls, device: int, base: int) -> None:
        with cls._lock:
            handle = cls._by_base.get((device, base))
            if handle is None:
                return
            ent = cls._by_handle[(device, handle)]
            ent[1] -= 1
            if ent[1] > 0:
                return
            del cls._by_handle[(device, handle)]
            del cls._by_base[(device, base)]
        ops.rt_ipc_close(device, base)


# ------------------------------------------------------------------ test harness
def emulate_exchange(
    models: Sequence[torch.Tensor], samples: Sequence[int], tile: int = 16384, wire_bf16: bool = True, seq: int = 1,
    timeout_cycles: int = 4_000_000_000,
) -> Tuple[List[torch.Tensor], torch.Tensor]:  # fmt: skip
    """All-to-all push + FedAvg among ``len(models)`` logical ranks on ONE GPU, in a
    single cooperative launch (B200_PROFILING.md: never co-schedule kernels that wait
    on one another as separate launches on one device)."""
    r = len(models)
    n = models[0].numel()
    dev = models[0].device.index or 0
    arenas = [DeviceArena(n, slots=r, tile=tile, wire_bf16=wire_bf16, device=dev) for _ in range(r)]
    try:
        bank = seq & 1
        total = float(sum(samples))
        pushes, reduces = [], []
        for me in range(r):
            arenas[me].params.copy_(models[me])
            peers = [p for p in range(r) if p != me]
            pushes.append(
                ops.make_push_table(
                    [arenas[p].slot_ptr(me, bank) for p in peers], [arenas[p].flag_ptr(me, bank) for p in peers], wire_bf16
                )
            )
            # contributors in rank order; the local model sits at position `me`
            reduces.append(
                ops.make_reduce_table(
                    [arenas[me].slot_ptr(p, bank) for p in peers],
                    [arenas[me].flag_ptr(p, bank) for p in peers],
                    [samples[p] / total for p in peers],
                    me,
                    samples[me] / total,
                    wire_bf16,
                    True,
                )
            )
        device = models[0].device
        push_dev = ops.tables_to_device(pushes, device)
        red_dev = ops.tables_to_device(reduces, device)
        status = torch.zeros(4 * r, dtype=torch.int32, device=device)
        ops.gossip_fused_emulated([a.params for a in arenas], tile, seq, push_dev, red_dev, timeout_cycles, status)
        torch.cuda.synchronize(dev)
        outs = [a.params.clone() for a in arenas]
        return outs, status.view(r, 4).cpu()
    finally:
        for a in arenas:
            a.close()
