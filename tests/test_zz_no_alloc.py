"""The allocation-free co-located exchange, driven through the C ABI.

Who needs it: a host program that cannot obtain device memory at the moment a round's
exchange is due - because it handed the whole GPU to its own pool when it started, or
because its round loop cannot absorb the device-wide synchronisation that ``cudaMalloc``
and ``cudaFree`` imply.  ``p2pfl_gossip_exchange_static`` keeps the inbox, the arrival
flags and the kernel tables in an array inside the library, so the caller passes only
the models.

Checked here, on two consecutive calls with different inputs (the second call meets the
scratch exactly as the first one left it): the result equals what the existing entry
point (``p2pfl_gossip_fused_emulated`` with caller-allocated arenas) produces; sizes
beyond the built-in limit are refused with a message; device memory in use does not grow.

This file sorts last so that a run which stops at the first failure has already run
every other test.
"""

import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _static_exchange(lib, models, weights, timeout_cycles=4_000_000_000, status=None):
    r = len(models)
    ptrs = (C.c_void_p * r)(*[m.data_ptr() for m in models])
    ws = (C.c_float * r)(*weights)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return lib.p2pfl_gossip_exchange_static(
        C.c_int(0), stream, C.c_int(r), ptrs, ws, C.c_uint64(models[0].numel()), C.c_uint64(timeout_cycles),
        C.c_void_p(0 if status is None else status.data_ptr()),
    )  # fmt: skip


def test_static_exchange_matches_the_allocating_entry_point_twice():
    from p2pfl_b200 import ops
    from p2pfl_b200.parallel.arena import emulate_exchange

    lib = ops.lib()
    lib.p2pfl_last_error.restype = C.c_char_p
    n, ranks = 235168, 4
    torch.cuda.set_device(0)
    for call in range(2):
        torch.manual_seed(100 + call)
        models = [torch.randn(n, device=DEV) * (call + 1) for _ in range(ranks)]
        samples = [100 + 37 * r + 11 * call for r in range(ranks)]
        total = float(sum(samples))
        weights = [s / total for s in samples]
        # existing entry point: inboxes allocated by the caller (DeviceArena -> cudaMalloc)
        expect, st = emulate_exchange([m.clone() for m in models], samples, tile=16384, wire_bf16=True, seq=1)
        assert st[:, 0].sum().item() == 0
        # new entry point: no inbox passed, nothing allocated
        mine = [m.clone() for m in models]
        status = torch.full((4 * ranks,), -1, dtype=torch.int32, device=DEV)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(0)[0]
        rc = _static_exchange(lib, mine, weights, status=status)
        assert rc == 0, lib.p2pfl_last_error()
        torch.cuda.synchronize()
        free_now = torch.cuda.mem_get_info(0)[0]
        if call > 0:  # the very first use may load the kernels lazily
            assert free_now >= free_before, "the second call consumed device memory"
        st2 = status.view(ranks, 4).cpu()
        assert st2[:, 0].sum().item() == 0, "a rank timed out"
        assert (st2[:, 1] > 0).all(), "status words were not written"
        # and against plain PyTorch: the fp32 weighted sum of the models as they travel (bf16 wire)
        oracle = sum(w * m.bfloat16().float() for w, m in zip(weights, models))
        for r in range(ranks):
            assert torch.allclose(mine[r], expect[r], atol=1e-5), f"call {call} rank {r}: {(mine[r] - expect[r]).abs().max().item()}"
            assert torch.allclose(mine[r], oracle, atol=2e-6 * (call + 1), rtol=1e-5), f"call {call} rank {r} vs fp32 oracle"
            assert torch.equal(mine[r], mine[0]), "ranks disagree bit-wise"


def test_static_exchange_at_its_limits():
    """The largest case the built-in scratch holds: 8 ranks x 262144 elements (a whole number of
    tiles), and 8 ranks with a partial last tile - both against the fp32 PyTorch sum."""
    from p2pfl_b200 import ops

    lib = ops.lib()
    lib.p2pfl_last_error.restype = C.c_char_p
    torch.cuda.set_device(0)
    for n in (262144, 262144 - 16384 + 8):
        torch.manual_seed(n)
        models = [torch.randn(n, device=DEV) for _ in range(8)]
        weights = [(r + 1) / 36.0 for r in range(8)]
        mine = [m.clone() for m in models]
        status = torch.full((32,), -1, dtype=torch.int32, device=DEV)
        assert _static_exchange(lib, mine, weights, status=status) == 0, lib.p2pfl_last_error()
        torch.cuda.synchronize()
        assert status.view(8, 4)[:, 0].sum().item() == 0
        oracle = sum(w * m.bfloat16().float() for w, m in zip(weights, models))
        for r in range(8):
            assert torch.allclose(mine[r], oracle, atol=2e-6, rtol=1e-5), (n, r)
            assert torch.equal(mine[r], mine[0])


def test_static_exchange_refuses_what_does_not_fit():
    from p2pfl_b200 import ops

    lib = ops.lib()
    lib.p2pfl_last_error.restype = C.c_char_p
    too_big = [torch.zeros(262144 + 8, device=DEV) for _ in range(2)]
    assert _static_exchange(lib, too_big, [0.5, 0.5]) != 0
    assert b"262144" in lib.p2pfl_last_error()
    small = [torch.zeros(4096, device=DEV) for _ in range(9)]
    assert _static_exchange(lib, small, [1 / 9] * 9) != 0
    assert b"ranks" in lib.p2pfl_last_error()
    # and a valid call afterwards still works
    ok = [torch.full((4096,), float(i), device=DEV) for i in range(2)]
    assert _static_exchange(lib, ok, [0.25, 0.75]) == 0
    torch.cuda.synchronize()
    assert torch.allclose(ok[0], torch.full((4096,), 0.75, device=DEV)) and torch.equal(ok[0], ok[1])
