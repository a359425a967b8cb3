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
def test_colocated_nodes_use_the_pull_path():
    from p2pfl_b200.communication.nvlink import NvlinkMemoryProtocol

    stats = _run_federation([DEV], NvlinkMemoryProtocol, rounds=2, n=3, train_set=2)
    assert all(s["host"] == 0 for s in stats), stats  # nothing went through pickle
    assert sum(s["pulled"] for s in stats) > 0
    assert sum(s["pulled_ms"] for s in stats) > 0.0  # device-timed (CUDA events on the copy stream)
    assert all(s["fused"] == 0 for s in stats)  # co-located: no kernels that wait on one another


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
