"""The NVLink data-plane protocol on real devices."""

import pytest
import torch

DEV = "cuda:0"


def _run_federation(devices, protocol, rounds=2, train_set=None, n=None):
    from p2pfl_b200.learning.b200_learner import B200Learner
    from p2pfl_b200.learning.data import MnistFederatedDM
    from p2pfl_b200.models import MLP
    from p2pfl_b200.node import Node
    from p2pfl_b200.settings import Settings
    from p2pfl_b200.utils import check_equal_models, wait_4_results, wait_convergence

    n = n or len(devices)
    if train_set:
        Settings.TRAIN_SET_SIZE = train_set
    nodes = [
        Node(MLP(seed=i), MnistFederatedDM(i, n, n_train=640 * n, n_test=128 * n), protocol=protocol,
             learner=B200Learner, device=devices[i % len(devices)])
        for i in range(n)
    ]  # fmt: skip
    for x in nodes:
        x.start()
    try:
        for i in range(1, n):
            for j in range(i):
                nodes[i].connect(nodes[j].addr)
        wait_convergence(nodes, n - 1, only_direct=True)
        nodes[0].set_start_learning(rounds=rounds, epochs=1)
        wait_4_results(nodes, 180)
        check_equal_models(nodes, atol=0)  # bit-identical on every node
        stats = [dict(x._communication_protocol.stats) for x in nodes]
        return stats
    finally:
        for x in nodes:
            x.stop()


@pytest.mark.gpu
def test_colocated_trainers_exchange_in_one_group_launch():
    """Several trainers of one process on ONE GPU (BASELINE config 4): their exchanges run as a
    single cooperative launch, never as kernels that wait on one another."""
    from p2pfl_b200.communication.nvlink import NvlinkMemoryProtocol

    stats = _run_federation([DEV], NvlinkMemoryProtocol, rounds=3, n=4, train_set=4)
    assert all(s["host"] == 0 for s in stats), stats  # nothing went through pickle
    assert all(s["fused"] == 3 for s in stats), stats
    assert all(s.get("fused_incomplete", 0) == 0 for s in stats), stats


@pytest.mark.gpu
def test_colocated_trainers_and_a_waiting_node():
    """Train set smaller than the network, all on one GPU: the trainers use the group launch, the
    node that waits gets the aggregate by pulling it."""
    from p2pfl_b200.communication.nvlink import NvlinkMemoryProtocol

    stats = _run_federation([DEV], NvlinkMemoryProtocol, rounds=2, n=3, train_set=2)
    assert all(s["host"] == 0 for s in stats), stats
    assert sum(s["fused"] for s in stats) == 4, stats  # two trainers x two rounds
    assert sum(s["pulled"] for s in stats) > 0
    assert sum(s["pulled_ms"] for s in stats) > 0.0  # device-timed (CUDA events on the copy stream)


def _group_exchange(models, samples, drop=None, timeout_cycles=0):
    """R members through p2pfl_gossip_fused_group on one GPU; ``drop``: a rank whose tables are left
    out of the launch (it never pushes), so everybody else misses it."""
    from p2pfl_b200 import ops
    from p2pfl_b200.parallel.arena import DeviceArena

    r, n = len(models), models[0].numel()
    dev = models[0].device
    arenas = [DeviceArena(n, slots=r, tile=16384, wire_bf16=True, device=dev.index or 0) for _ in range(r)]
    snaps = [torch.full_like(m, float("nan")) for m in models]
    try:
        total = float(sum(samples))
        members, pushes, reds = [], [], []
        live = [k for k in range(r) if k != drop]
        for me in live:
            arenas[me].params.copy_(models[me])
            peers = [p for p in range(r) if p != me]
            pushes.append(ops.make_push_table([arenas[p].slot_ptr(me, 1) for p in peers], [arenas[p].flag_ptr(me, 1) for p in peers], True))
            reds.append(ops.make_reduce_table([arenas[me].slot_ptr(p, 1) for p in peers], [arenas[me].flag_ptr(p, 1) for p in peers],
                                              [samples[p] / total for p in peers], me, samples[me] / total, True, True))  # fmt: skip
            members.append(ops.make_group_member(arenas[me].params, None, snaps[me], arenas[me].abort, arenas[me].status))
        md, pd, rd = (ops.tables_to_device(t, dev) for t in (members, pushes, reds))
        ops.gossip_fused_group(dev.index or 0, n, 16384, 1, md, pd, rd, len(live), timeout_cycles)
        torch.cuda.synchronize(dev)
        return ([arenas[k].params.clone() for k in range(r)], [arenas[k].status.cpu() for k in range(r)], snaps)
    finally:
        for a in arenas:
            a.close()


@pytest.mark.gpu
def test_group_kernel_matches_fp32_average_and_keeps_a_snapshot():
    torch.manual_seed(3)
    n, r = 235264, 5
    models = [torch.randn(n, device=DEV) for _ in range(r)]
    samples = [100, 200, 300, 150, 250]
    outs, status, snaps = _group_exchange(models, samples)
    ref = sum(w / sum(samples) * m.bfloat16().float() for w, m in zip(samples, models))
    for k in range(r):
        assert int(status[k][0]) == 0 and int(status[k][2]) == 0
        assert torch.equal(outs[k], outs[0])  # identical bits on every member
        assert torch.allclose(outs[k], ref, atol=2e-6, rtol=1e-5)
        assert torch.equal(snaps[k], models[k])  # the pre-exchange model, exactly


@pytest.mark.gpu
def test_group_kernel_reports_who_was_missing():
    """A member that never pushes: the others time out on its tiles, say so in the status words
    (flag + position mask) and still hold their own model in the snapshot."""
    torch.manual_seed(4)
    n, r = 65536, 3
    models = [torch.randn(n, device=DEV) for _ in range(r)]
    outs, status, snaps = _group_exchange(models, [1, 1, 1], drop=2, timeout_cycles=2_000_000)
    for k in (0, 1):
        assert int(status[k][0]) == 1
        assert int(status[k][2]) == 0b10  # table position 1 of ranks 0 and 1 is rank 2
        assert torch.equal(snaps[k], models[k])


@pytest.mark.gpu
@pytest.mark.multigpu
def test_fused_exchange_across_gpus_in_process():
    from p2pfl_b200.communication.nvlink import NvlinkMemoryProtocol

    ng = min(torch.cuda.device_count(), 4)
    stats = _run_federation([f"cuda:{i}" for i in range(ng)], NvlinkMemoryProtocol, rounds=3, train_set=ng)
    assert all(s["host"] == 0 for s in stats), stats
    assert all(s["fused"] == 3 for s in stats), stats
    assert all(s["fused_ms"] > 0.0 for s in stats), stats  # device-timed, includes the wait for peers


@pytest.mark.gpu
@pytest.mark.multigpu
def test_fused_exchange_with_non_trainers():
    """Train set smaller than the network: trainers run the fused kernel, the others pull."""
    from p2pfl_b200.communication.nvlink import NvlinkMemoryProtocol

    ng = torch.cuda.device_count()
    stats = _run_federation([f"cuda:{i}" for i in range(ng)], NvlinkMemoryProtocol, rounds=2, train_set=2, n=ng + 1)
    assert all(s["host"] == 0 for s in stats), stats
    assert sum(s["pulled"] for s in stats) > 0


@pytest.mark.gpu
@pytest.mark.multigpu
def test_fused_kernel_dead_peer_is_aborted_from_the_host():
    """The real (one launch per GPU) fused kernel with a peer on ANOTHER GPU that never pushes: the
    kernel has pushed its own tiles and spins on the peer's flags; the host flips the abort word,
    the kernel ends, reports who was missing, and the snapshot still holds the pre-exchange model."""
    import time

    from p2pfl_b200 import ops
    from p2pfl_b200.parallel.arena import DeviceArena

    n = 4 * 16384
    a = DeviceArena(n, slots=2, tile=16384, wire_bf16=True, device=0)
    b = DeviceArena(n, slots=2, tile=16384, wire_bf16=True, device=1)  # the peer that stays silent
    try:
        ops.rt_enable_peer(0, 1)
        torch.manual_seed(5)
        model = torch.randn(n, device="cuda:0")
        a.params.copy_(model)
        snap = torch.full_like(model, float("nan"))
        push = ops.make_push_table([b.slot_ptr(0, 1)], [b.flag_ptr(0, 1)], True)
        red = ops.make_reduce_table([a.slot_ptr(1, 1)], [a.flag_ptr(1, 1)], [0.5], 0, 0.5, True, True)
        side = torch.cuda.Stream(device="cuda:0")
        with torch.cuda.device(0):
            a.status.zero_()
            a.abort.zero_()
            torch.cuda.synchronize(0)
            with torch.cuda.stream(side):
                ops.gossip_fused(a.params, None, 16384, 1, push, red, a.abort, 0, a.status, snapshot=snap)  # no device timeout
                done = torch.cuda.Event()
                done.record(side)
            time.sleep(0.05)
            assert not done.query(), "the kernel must still be waiting for the silent peer"
            a.abort.fill_(1)  # (default stream: the kernel runs on `side`)
            t0 = time.time()
            while not done.query() and time.time() - t0 < 10:
                time.sleep(0.001)
            assert done.query(), "the abort word did not release the kernel"
            torch.cuda.synchronize(0)
        st = a.status.cpu()
        assert int(st[0]) == 1 and int(st[2]) == 1  # aborted; table position 0 (the peer) was missing
        assert torch.equal(snap, model)  # what the host puts back before it falls back to add_model
        # our own tiles did reach the peer's inbox before the wait
        torch.cuda.synchronize(1)
        assert torch.equal(b.slot_tensor(0, 1).to("cuda:0").float(), model.bfloat16().float())
    finally:
        a.close()
        b.close()
