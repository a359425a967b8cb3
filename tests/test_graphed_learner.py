"""GraphedTorchLearner: PyTorch numerics, one CUDA-graph replay per step (learning/graphed_learner.py)."""

import pytest
import torch
import torch.nn.functional as F

from p2pfl_b200.learning.data import MnistFederatedDM
from p2pfl_b200.learning.graphed_learner import GraphedTorchLearner
from p2pfl_b200.models import MLP


def _dm():
    return MnistFederatedDM(0, 1, batch_size=32, n_train=1100, n_test=200, seed=5)


def test_without_cuda_it_is_the_eager_learner():
    lr = GraphedTorchLearner(MLP(seed=0), _dm(), "n0", epochs=1, device="cpu")
    assert not lr.graphed and lr.stream is None
    before = lr.evaluate()["test_loss"]
    lr.fit()
    after = lr.evaluate()
    assert after["test_loss"] < before and lr.h2d_bytes == 0


@pytest.mark.gpu
def test_graphed_steps_equal_eager_steps():
    """Same batches in the same order: the replayed graph and a plain eager loop (ordinary Adam) end
    with the same parameters; the short last batch, the per-round H2D accounting and a second round
    (fresh Adam state, as the reference's fresh Trainer) are covered."""
    from p2pfl_b200.settings import Settings

    torch.backends.cuda.matmul.allow_tf32 = False
    dm = _dm()
    lr = GraphedTorchLearner(MLP(seed=3), dm, "n0", epochs=1, device="cuda:0", shuffle=False)
    assert lr.graphed
    ref = MLP(seed=3).to("cuda:0")
    x, y = (t.to("cuda:0") for t in dm.tensors("train"))
    n = x.shape[0]
    assert n % 32 != 0  # there is a short last batch
    for rnd in range(2):
        lr.fit()
        opt = torch.optim.Adam(ref.parameters(), lr=float(getattr(ref, "lr", 1e-3)))
        assert Settings.RESET_OPTIMIZER_EACH_ROUND
        ref.train()
        for i in range(0, n, 32):
            opt.zero_grad(set_to_none=True)
            F.cross_entropy(ref(x[i : i + 32]), y[i : i + 32]).backward()
            opt.step()
        torch.cuda.synchronize()
        for (name, p), q in zip(lr.model.named_parameters(), ref.parameters()):
            # (capturable foreach Adam and the ordinary one order their arithmetic differently: ~2e-5 after two rounds)
            assert torch.allclose(p, q, atol=5e-4, rtol=1e-3), (rnd, name, (p - q).abs().max().item())
    per_round = sum(t.numel() * t.element_size() for s in ("train", "val") for t in dm.tensors(s))
    assert lr.h2d_bytes == 2 * per_round
    res = lr.evaluate()
    ref.eval()
    xt, yt = (t.to("cuda:0") for t in dm.tensors("test"))
    with torch.no_grad():
        want = F.cross_entropy(ref(xt), yt).item()
    assert abs(res["test_loss"] - want) < 1e-2 * max(1.0, want) and 0.0 <= res["test_metric"] <= 1.0


@pytest.mark.gpu
def test_graphed_learner_in_a_node():
    """Two nodes, in-memory protocol, graphed learners on the GPU: a full experiment converges to equal models."""
    from p2pfl_b200.communication.memory import InMemoryCommunicationProtocol
    from p2pfl_b200.node import Node
    from p2pfl_b200.utils import check_equal_models, wait_4_results, wait_convergence

    nodes = [
        Node(MLP(seed=i), MnistFederatedDM(i, 2, batch_size=32, n_train=1600, n_test=200), learner=GraphedTorchLearner,
             protocol=InMemoryCommunicationProtocol, device="cuda:0")
        for i in range(2)
    ]  # fmt: skip
    try:
        for nd in nodes:
            nd.start()
        nodes[1].connect(nodes[0].addr)
        wait_convergence(nodes, 1, only_direct=True)
        nodes[0].set_start_learning(rounds=2, epochs=1)
        wait_4_results(nodes)
        check_equal_models(nodes)
    finally:
        for nd in nodes:
            nd.stop()
