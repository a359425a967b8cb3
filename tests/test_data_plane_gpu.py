"""Copy-engine (cp.async.bulk) and NVLS (multimem) variants of the data plane against plain PyTorch."""

import pytest
import torch

pytestmark = pytest.mark.gpu
TILE = 16384


def _arena(n, dev=0, slots=2, bf16=True):
    from p2pfl_b200.parallel.arena import DeviceArena

    return DeviceArena(n, slots=slots, tile=TILE, wire_bf16=bf16, device=dev)


def _devices():
    return [0, 1] if torch.cuda.device_count() >= 2 else [0, 0]


@pytest.mark.parametrize("bf16", [True, False])
def test_bulk_push_into_an_inbox_slot(bf16):
    """One bulk load + one bulk store per tile and peer; flags take the sequence number."""
    from p2pfl_b200 import ops

    d0, d1 = _devices()
    n = 5 * TILE + 4096  # a partial last tile
    a, b = _arena(n, d0, bf16=bf16), _arena(n, d1, bf16=bf16)
    try:
        if d0 != d1:
            ops.rt_enable_peer(d0, d1)
        torch.manual_seed(0)
        src = torch.randn(n, device=f"cuda:{d0}")
        a.params.copy_(src)
        b.slot_tensor(1, 0).fill_(float("nan"))
        table = ops.make_push_table([b.slot_ptr(1, 0)], [b.flag_ptr(1, 0)], bf16)
        with torch.cuda.device(d0):
            ops.gossip_push_bulk(a.params, 7, table)
            torch.cuda.synchronize(d0)
        got = b.slot_tensor(1, 0).float().to(f"cuda:{d0}")
        want = src.bfloat16().float() if bf16 else src
        assert torch.equal(got, want)
        assert torch.equal(b.flags_tensor(1, 0).cpu(), torch.full((6,), 7, dtype=torch.int32))
        assert int(b.flags_tensor(0, 0).abs().sum()) == 0
    finally:
        a.close()
        b.close()


def test_bulk_store_of_a_sub_range_into_peer_parameters():
    """G3 (init / diffusion) with an (offset, length) range: only the named tiles move."""
    from p2pfl_b200 import ops

    d0, d1 = _devices()
    n = 6 * TILE
    a, b = _arena(n, d0, bf16=False), _arena(n, d1, bf16=False)
    try:
        if d0 != d1:
            ops.rt_enable_peer(d0, d1)
        src = torch.arange(n, dtype=torch.float32, device=f"cuda:{d0}")
        a.params.copy_(src)
        b.params.fill_(-1.0)
        table = ops.make_push_table([b.params.data_ptr()], [b.flag_ptr(0, 1)], False)
        with torch.cuda.device(d0):
            ops.gossip_push_bulk(a.params, 3, table, begin=2 * TILE, end=5 * TILE)
            torch.cuda.synchronize(d0)
        got = b.params.to(f"cuda:{d0}")
        assert torch.equal(got[2 * TILE : 5 * TILE], src[2 * TILE : 5 * TILE])
        assert float(got[: 2 * TILE].max()) == -1.0 and float(got[5 * TILE :].max()) == -1.0
        assert b.flags_tensor(0, 1).cpu().tolist() == [0, 0, 3, 3, 3, 0]
    finally:
        a.close()
        b.close()


def test_remote_reduce_mode_matches_fp32_average():
    """Mode A: two senders' copy engines add w_i * model_i into the receiver's accumulator; the
    receiver only finalises.  Compared with the fp32 weighted sum (addition order is free: tolerance)."""
    from p2pfl_b200 import ops

    d0, d1 = _devices()
    n = 3 * TILE + 8192
    arenas = [_arena(n, d0, slots=3, bf16=False), _arena(n, d1, slots=3, bf16=False), _arena(n, d0, slots=3, bf16=False)]
    recv = arenas[2]
    try:
        if d0 != d1:
            ops.rt_enable_peer(d1, d0)
        torch.manual_seed(1)
        models = [torch.randn(n, device=a.params.device) for a in arenas]
        w = [0.2, 0.5, 0.3]
        for a, m in zip(arenas, models):
            a.params.copy_(m)
        recv.outbox.zero_()  # the accumulator
        counts = recv.flags_tensor(0, 0)
        counts.zero_()
        for i in (0, 1):
            table = ops.make_push_table([recv.outbox.data_ptr()], [counts.data_ptr()], False)
            with torch.cuda.device(arenas[i].device):
                ops.gossip_push_reduce_bulk(arenas[i].params, w[i], table)
        status = torch.zeros(4, dtype=torch.int32, device=recv.params.device)
        w16 = torch.empty(n, dtype=torch.bfloat16, device=recv.params.device)
        with torch.cuda.device(recv.device):
            ops.gossip_finalize_acc(recv.params, w16, recv.outbox, counts, 2, w[2], int(4e9), status)
        for d in {d0, d1}:
            torch.cuda.synchronize(d)
        ref = sum(wi * m.to(recv.params.device) for wi, m in zip(w, models))
        assert int(status[0]) == 0 and int(status[1]) == 4
        assert torch.allclose(recv.params, ref, atol=1e-6, rtol=1e-6)
        assert torch.equal(w16, recv.params.bfloat16())
        assert float(recv.outbox.abs().max()) == 0.0 and int(counts.abs().sum()) == 0  # ready for the next exchange
    finally:
        for a in arenas:
            a.close()


def test_multicast_allreduce_and_store_or_clean_fallback():
    """NVLS path when the device offers it (in-process group over the visible GPUs), else the probe
    must say no without raising - callers then use the per-peer kernels."""
    from p2pfl_b200.parallel import multicast as mc

    if not mc.supported(0):
        pytest.skip("multicast not supported on this device (probe returned 0, callers fall back)")
    devs = list(range(min(torch.cuda.device_count(), 4)))
    n = 2 * TILE + 1024
    try:
        group = mc.MulticastGroup.in_process(devs, n)
    except RuntimeError as e:
        pytest.skip(f"multicast object could not be created here: {e}")
    try:
        torch.manual_seed(2)
        models = [torch.randn(n, device=f"cuda:{d}") for d in devs]
        w = [1.0 / len(devs)] * len(devs)
        params = [m.clone() for m in models]
        w16 = [torch.empty(n, dtype=torch.bfloat16, device=m.device) for m in models]
        for r, d in enumerate(devs):
            with torch.cuda.device(d):
                group.bindings[r].allreduce(params[r], w16[r], r, len(devs), w[r], int(8e9))
        for d in devs:
            torch.cuda.synchronize(d)
        ref = sum(wi * m.to("cuda:0") for wi, m in zip(w, models))
        for r in range(len(devs)):
            assert torch.allclose(params[r].to("cuda:0"), ref, atol=1e-6, rtol=1e-6)
            assert torch.equal(w16[r], params[r].bfloat16())
        # one multimem.st reaches every rank's copy
        with torch.cuda.device(devs[0]):
            group.bindings[0].store(models[0], 5)
            torch.cuda.synchronize(devs[0])
        for r in range(len(devs)):
            assert torch.equal(group.bindings[r].local.to("cuda:0"), models[0].to("cuda:0"))
            assert int(group.bindings[r].local_flags[2]) == 5
    finally:
        group.close()
