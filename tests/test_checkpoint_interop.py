"""Checkpoint/resume, and wire interoperability with the UNMODIFIED reference (if installed)."""

import os
import sys
import time

import pytest
import torch

from p2pfl_b200.communication.grpc import GrpcCommunicationProtocol
from p2pfl_b200.communication.memory import InMemoryCommunicationProtocol
from p2pfl_b200.learning.checkpoint import load_checkpoint, save_checkpoint
from p2pfl_b200.learning.data import MnistFederatedDM
from p2pfl_b200.models import MLP
from p2pfl_b200.node import Node
from p2pfl_b200.utils import wait_4_results, wait_convergence

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")


def test_checkpoint_roundtrip(tmp_path):
    dm = MnistFederatedDM(0, 1, n_train=320, n_test=64)
    n = Node(MLP(seed=5), dm, protocol=InMemoryCommunicationProtocol, device="cpu")
    n.start()
    try:
        n.set_start_learning(rounds=1, epochs=1)
        wait_4_results([n], 60)
        trained = {k: v.clone() for k, v in n.state.learner.get_parameters().items()}
        path = str(tmp_path / "ckpt.pt")
        save_checkpoint(n, path)
    finally:
        n.stop()
    fresh = Node(MLP(seed=9), dm, protocol=InMemoryCommunicationProtocol, device="cpu")
    ck = load_checkpoint(fresh, path)
    assert ck["total_rounds"] is None or ck["total_rounds"] >= 0
    for k, v in fresh.model.state_dict().items():
        assert torch.equal(v, trained[k])
    with pytest.raises(ValueError):
        from p2pfl_b200.models import CNN

        load_checkpoint(Node(CNN(), dm, protocol=InMemoryCommunicationProtocol, device="cpu"), path)


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "p2pfl")), reason="reference not installed (baseline/install_reference.sh)")
def test_interop_with_reference_node_over_grpc():
    """A reference node and one of ours handshake over gRPC, see each other's heartbeats,
    and run a protocol-only experiment (epochs=0) started by the REFERENCE node: its pickled
    initial model must load here byte-compatibly and ours must be accepted by its aggregator."""
    sys.path[:0] = [os.path.join(ROOT, "baseline", "shim"), REF]
    try:
        from p2pfl.learning.pytorch.mnist_examples.models.mlp import MLP as RefMLP
        from p2pfl.node import Node as RefNode
        from p2pfl.settings import Settings as RefSettings
    except Exception as e:  # pragma: no cover
        pytest.skip(f"reference not importable: {e}")
    RefSettings.GRPC_TIMEOUT = 2
    RefSettings.HEARTBEAT_PERIOD = 0.5
    RefSettings.HEARTBEAT_TIMEOUT = 5
    RefSettings.WAIT_HEARTBEATS_CONVERGENCE = 0.4
    RefSettings.GOSSIP_MODELS_PERIOD = 0.2
    RefSettings.GOSSIP_PERIOD = 0
    RefSettings.AGGREGATION_TIMEOUT = 20
    RefSettings.VOTE_TIMEOUT = 20
    dm = MnistFederatedDM(0, 2, n_train=640, n_test=128)
    ref = RefNode(RefMLP(), dm)
    mine = Node(MLP(seed=3), MnistFederatedDM(1, 2, n_train=640, n_test=128), protocol=GrpcCommunicationProtocol, device="cpu")
    ref.start()
    mine.start()
    try:
        assert mine.connect(ref.addr)
        wait_convergence([mine], 1, only_direct=True, wait=10)
        deadline = time.time() + 10
        while len(ref.get_neighbors(only_direct=True)) != 1 and time.time() < deadline:
            time.sleep(0.05)
        assert len(ref.get_neighbors(only_direct=True)) == 1
        init = {k: v.clone() for k, v in ref.learner_class and ref.model.state_dict().items()}
        ref.set_start_learning(rounds=1, epochs=0)
        wait_4_results([mine], 90)
        deadline = time.time() + 60
        while ref.state.round is not None and time.time() < deadline:
            time.sleep(0.1)
        assert ref.state.round is None, "reference node did not finish"
        got = mine.state.learner.get_parameters()
        for k, v in init.items():
            assert torch.allclose(got[k].cpu(), v, atol=1e-6), k
        # still neighbours: nobody evicted anybody over a malformed message
        assert len(mine.get_neighbors(only_direct=True)) == 1
    finally:
        try:
            ref.stop()
        finally:
            mine.stop()
