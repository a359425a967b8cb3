"""Unit tests of the learning layer (mirrors reference test/learning_test.py)."""

import pickle

import numpy as np
import pytest
import torch

from p2pfl_b200.exceptions import DecodingParamsError, ModelNotMatchingError, NoModelsToAggregateError
from p2pfl_b200.learning.aggregators.fedavg import FedAvg, fedavg_state_dicts
from p2pfl_b200.learning.data import MnistFederatedDM
from p2pfl_b200.learning.params import FlatParams, ParamLayout
from p2pfl_b200.learning.torch_learner import TorchLearner
from p2pfl_b200.models import CNN, MLP, ResNet18, gpt2_tiny


def _dm():
    return MnistFederatedDM(0, 1, n_train=320, n_test=64)


def test_encoding_roundtrip_is_byte_identical():
    l1 = TorchLearner(MLP(seed=1), _dm(), device="cpu")
    enc = l1.encode_parameters()
    l2 = TorchLearner(MLP(seed=2), _dm(), device="cpu")
    l2.set_parameters(l2.decode_parameters(enc))
    assert l2.encode_parameters() == enc


def test_wire_format_is_pickled_list_of_ndarrays_in_state_dict_order():
    model = MLP(seed=3)
    ref = [v.detach().numpy().copy() for v in model.state_dict().values()]
    learner = TorchLearner(model, _dm(), device="cpu")
    arrays = pickle.loads(learner.encode_parameters())
    assert isinstance(arrays, list) and len(arrays) == 6
    for a, r in zip(arrays, ref):
        assert a.dtype == np.float32 and np.array_equal(a, r)
    assert sum(a.size for a in arrays) == 235146


def test_decode_rejects_code_execution_and_garbage():
    learner = TorchLearner(MLP(), _dm(), device="cpu")

    class Evil:
        def __reduce__(self):
            return (eval, ("1+1",))

    with pytest.raises(DecodingParamsError):
        learner.decode_parameters(pickle.dumps([Evil()]))
    with pytest.raises(DecodingParamsError):
        learner.decode_parameters(b"not a pickle")
    with pytest.raises(DecodingParamsError):
        learner.decode_parameters(pickle.dumps({"a": 1}))


def test_decode_rejects_other_architecture():
    mlp = TorchLearner(MLP(), _dm(), device="cpu")
    cnn = TorchLearner(CNN(), _dm(), device="cpu")
    with pytest.raises(ModelNotMatchingError):
        mlp.decode_parameters(cnn.encode_parameters())
    # same tensor count, wrong shapes
    other = TorchLearner(MLP(hidden=(64, 32)), _dm(), device="cpu")
    with pytest.raises(ModelNotMatchingError):
        mlp.decode_parameters(other.encode_parameters())


def test_avg_simple():
    agg = FedAvg()
    a = {"a": torch.tensor(-1.0), "b": torch.tensor(-1.0)}
    b = {"a": torch.tensor(0.0), "b": torch.tensor(0.0)}
    c = {"a": torch.tensor(1.0), "b": torch.tensor(1.0)}
    r = agg.aggregate({"a": (a, 1), "b": (b, 1), "c": (c, 1)})
    assert all(float(v) == 0.0 for v in r.values())
    r = agg.aggregate({"a": (a, 1), "b": (b, 7), "c": (c, 1)})
    assert all(float(v) == 0.0 for v in r.values())
    r = agg.aggregate({"a": (a, 800), "b": (b, 0), "c": (c, 0)})
    assert all(float(v) == -1.0 for v in r.values())
    with pytest.raises(NoModelsToAggregateError):
        agg.aggregate({})


def test_avg_complex_flat_matches_state_dict_oracle():
    learner = TorchLearner(MLP(seed=0), _dm(), device="cpu")
    base = learner.export_model()
    assert isinstance(base, FlatParams)
    # identity
    same = FedAvg().aggregate({"x": (base, 3)})
    assert torch.equal(same.flat, base.flat)
    up = FlatParams(base.flat + 1, base.layout)
    down = FlatParams(base.flat - 1, base.layout)
    mid = FedAvg().aggregate({"u": (up, 1), "d": (down, 1)})
    assert torch.allclose(mid.flat, base.flat, atol=1e-6)
    # weighted, vs. the per-tensor oracle
    rnd = FlatParams(torch.randn_like(base.flat), base.layout)
    flat = FedAvg().aggregate({"a": (base, 5), "b": (rnd, 2), "c": (up, 9)})
    oracle = fedavg_state_dicts({"a": (base.state_dict(), 5), "b": (rnd.state_dict(), 2), "c": (up.state_dict(), 9)})
    for k, v in flat.state_dict().items():
        assert torch.allclose(v, oracle[k], atol=1e-6)


def test_layout_alignment_and_int_buffers():
    model = ResNet18()
    sd = model.state_dict()
    layout = ParamLayout.from_state_dict(sd)
    assert len(layout.slots) == 122
    assert sum(1 for s in layout.slots if not s.in_arena) == 20  # num_batches_tracked
    assert all(s.offset % 32 == 0 for s in layout.slots if s.in_arena)
    assert layout.num_params == sum(v.numel() for v in sd.values() if v.is_floating_point())
    learner = TorchLearner(model, None, device="cpu")
    enc = learner.encode_parameters()
    learner2 = TorchLearner(ResNet18(), None, device="cpu")
    learner2.set_parameters(learner2.decode_parameters(enc))
    assert learner2.encode_parameters() == enc
    avg = FedAvg().aggregate({"a": (learner.export_model(), 1), "b": (learner2.export_model(), 3)})
    assert avg.extra and all(not v.is_floating_point() for v in avg.extra.values())


def test_params_are_views_of_the_arena_and_training_moves_them():
    dm = _dm()
    learner = TorchLearner(MLP(seed=0), dm, device="cpu", epochs=1)
    before = learner.flat.clone()
    p = dict(learner.model.named_parameters())["l1.weight"]
    assert p.data_ptr() == learner.flat.data_ptr()  # first slot starts the arena
    r0 = learner.evaluate()
    learner.fit()
    r1 = learner.evaluate()
    assert not torch.equal(before, learner.flat)
    assert r1["test_loss"] < r0["test_loss"]
    assert learner.get_num_samples() == (288, 64)


def test_gpt2_tiny_learner_step():
    from p2pfl_b200.learning.data import TokenFederatedDM

    dm = TokenFederatedDM(0, 1, batch_size=4, seq_len=16, vocab=256, n_train=16, n_test=8)
    learner = TorchLearner(gpt2_tiny(seed=0), dm, device="cpu", epochs=1)
    r0 = learner.evaluate()
    learner.fit()
    r1 = learner.evaluate()
    assert r1["test_loss"] < r0["test_loss"]
