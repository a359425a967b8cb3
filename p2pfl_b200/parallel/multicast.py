"""NVLS multicast groups (NVSwitch multicast + in-switch reduction) for the model exchange.

A group spans N GPUs; every rank binds an allocation of its own to it and gets two views of the
same bytes: ``local`` (its own copy) and ``mc`` (the multicast address: a store reaches every
rank's copy, a ``multimem.ld_reduce`` returns the sum over the ranks).  Used for the multicast
variant of the model store (init / diffusion) and for the two-shot weighted all-reduce that bounds
config 5 from above.  ``supported()`` probes the device; callers fall back to the
per-peer kernels.  No counterpart in the reference.
"""

from __future__ import annotations

import ctypes as C
import os
import socket
import tempfile
from typing import Any, List, Optional

import torch

from p2pfl_b200 import ops
from p2pfl_b200.parallel.arena import tensor_from_ptr

FLAG_BYTES = 256


def supported(device: int = 0) -> bool:
    try:
        return bool(ops.lib().p2pfl_mc_supported(C.c_int(device)))
    except Exception:
        return False


def _rounded(n_devices: int, nbytes: int) -> int:
    out = C.c_uint64(0)
    ops._check(ops.lib().p2pfl_mc_rounded_size(C.c_int(n_devices), C.c_uint64(nbytes), C.byref(out)))
    return int(out.value)


class Binding:
    """One rank's share of a group: ``local`` / ``mc`` fp32 views of ``n`` elements + the flag words."""

    def __init__(self, group: C.c_void_p, device: int, n: int, size: int) -> None:
        self.device, self.n, self.size = device, n, size
        self.flags_offset = size - FLAG_BYTES
        b, lp, mp = C.c_void_p(), C.c_void_p(), C.c_void_p()
        ops._check(ops.lib().p2pfl_mc_bind(group, C.c_int(device), C.byref(b), C.byref(lp), C.byref(mp)))
        self._b, self.local_ptr, self.mc_ptr = b, int(lp.value), int(mp.value)
        self.local = tensor_from_ptr(self.local_ptr, n, torch.float32, device)
        self.local_flags = tensor_from_ptr(self.local_ptr + self.flags_offset, FLAG_BYTES // 4, torch.int32, device)
        self.epoch = 0

    def allreduce(self, params: torch.Tensor, params_bf16: Optional[torch.Tensor], rank: int, n_ranks: int, weight: float,
                  timeout_cycles: int = 0, status: Optional[torch.Tensor] = None) -> None:  # fmt: skip
        """params <- sum over ranks of weight_r * params_r, through the switch (every rank calls it)."""
        self.epoch += 1
        ops._check(
            ops.lib().p2pfl_mc_allreduce(
                C.c_int(self.device), ops._stream(), ops._p(params), ops._p(params_bf16), C.c_void_p(self.local_ptr),
                C.c_void_p(self.mc_ptr), C.c_uint64(params.numel()), C.c_uint64(self.flags_offset), C.c_int(rank),
                C.c_int(n_ranks), C.c_float(weight), C.c_uint32(self.epoch), C.c_uint64(timeout_cycles), ops._p(status),
            )
        )  # fmt: skip
        ops._count()

    def store(self, src: torch.Tensor, seq: int) -> None:
        """``src`` lands in every rank's ``local`` (one multimem.st per 16 bytes), then flag word 2 = seq."""
        ops._check(
            ops.lib().p2pfl_mc_store(C.c_int(self.device), ops._stream(), ops._p(src), C.c_void_p(self.mc_ptr),
                                     C.c_uint64(src.numel()), C.c_uint64(self.flags_offset + 8), C.c_uint32(seq))
        )  # fmt: skip
        ops._count()

    def close(self) -> None:
        if self._b is not None:
            self.local = self.local_flags = None  # type: ignore[assignment]
            ops.lib().p2pfl_mc_unbind(self._b)
            self._b = None


class MulticastGroup:
    def __init__(self, handle: C.c_void_p, n: int, size: int) -> None:
        self._h, self.n, self.size = handle, n, size
        self.bindings: List[Binding] = []

    @classmethod
    def in_process(cls, devices: List[int], n: int) -> "MulticastGroup":
        """All ranks are GPUs of this process."""
        size = _rounded(len(devices), n * 4 + FLAG_BYTES)
        h = C.c_void_p()
        ops._check(ops.lib().p2pfl_mc_create(C.c_int(devices[0]), C.c_int(len(devices)), C.c_uint64(size), C.byref(h), None))
        g = cls(h, n, size)
        for d in devices:
            ops._check(ops.lib().p2pfl_mc_add_device(h, C.c_int(d)))
        for d in devices:
            g.bindings.append(Binding(h, d, n, size))
        return g

    @classmethod
    def distributed(cls, dist: Any, device: int, n: int) -> "MulticastGroup":
        """One rank per process (``torch.distributed`` is initialised): rank 0 creates the object and hands
        its file descriptor to the others over a unix socket (SCM_RIGHTS)."""
        rank, world = dist.get_rank(), dist.get_world_size()
        size = _rounded(world, n * 4 + FLAG_BYTES)
        h = C.c_void_p()
        path = [os.path.join(tempfile.gettempdir(), f"p2pfl_mc_{os.getpid()}_{n}.sock") if rank == 0 else None]
        if rank == 0:
            fd = C.c_int(-1)
            ops._check(ops.lib().p2pfl_mc_create(C.c_int(device), C.c_int(world), C.c_uint64(size), C.byref(h), C.byref(fd)))
            if os.path.exists(path[0]):
                os.unlink(path[0])
            srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            srv.bind(path[0])
            srv.listen(world)
        dist.broadcast_object_list(path, src=0)
        if rank == 0:
            for _ in range(world - 1):
                conn, _ = srv.accept()
                socket.send_fds(conn, [b"mc"], [fd.value])
                conn.close()
            srv.close()
            os.unlink(path[0])
            os.close(fd.value)
        else:
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            c.connect(path[0])
            _, fds, _, _ = socket.recv_fds(c, 16, 1)
            c.close()
            ops._check(ops.lib().p2pfl_mc_import(C.c_int(device), C.c_int(fds[0]), C.c_int(world), C.c_uint64(size), C.byref(h)))
            os.close(fds[0])
        ops._check(ops.lib().p2pfl_mc_add_device(h, C.c_int(device)))
        dist.barrier()  # every device is part of the object before anybody binds memory to it
        g = cls(h, n, size)
        g.bindings.append(Binding(h, device, n, size))
        dist.barrier()
        return g

    def close(self) -> None:
        for b in self.bindings:
            b.close()
        self.bindings = []
        if self._h is not None:
            ops.lib().p2pfl_mc_destroy(self._h)
            self._h = None
