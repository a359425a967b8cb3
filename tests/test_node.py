"""End-to-end learning tests with real nodes (mirrors reference test/node_test.py
and adds what the reference leaves untested: in-memory transport, epochs > 0,
non-trainers, vote tally, aggregator admission, re-running an experiment)."""

import time

import pytest
import torch

from p2pfl_b200.communication.grpc import GrpcCommunicationProtocol
from p2pfl_b200.communication.memory import InMemoryCommunicationProtocol
from p2pfl_b200.exceptions import NodeRunningException, ZeroRoundsException
from p2pfl_b200.learning.aggregators.fedavg import FedAvg
from p2pfl_b200.learning.data import MnistFederatedDM
from p2pfl_b200.management.logger import logger
from p2pfl_b200.models import CNN, MLP
from p2pfl_b200.node import Node
from p2pfl_b200.settings import Settings
from p2pfl_b200.stages.base_node.vote_train_set_stage import tally_votes
from p2pfl_b200.utils import check_equal_models, wait_4_results, wait_convergence

from p2pfl_b200.communication.sockets import SocketCommunicationProtocol  # noqa: E402

PROTOCOLS = [InMemoryCommunicationProtocol, GrpcCommunicationProtocol, SocketCommunicationProtocol]


def _dm(i=0, n=1):
    return MnistFederatedDM(i, n, n_train=640 * n, n_test=64 * n)


def _line(n, protocol, model=MLP, device="cpu"):
    nodes = [Node(model(), _dm(i, n), protocol=protocol, device=device) for i in range(n)]
    for x in nodes:
        x.start()
    for i in range(n - 1):
        nodes[i + 1].connect(nodes[i].addr)
    wait_convergence(nodes, n - 1, only_direct=False)
    return nodes


def _stop(nodes):
    for x in nodes:
        x.stop()


@pytest.mark.parametrize("protocol", PROTOCOLS)
@pytest.mark.parametrize("n,r", [(2, 1), (2, 2)])
def test_convergence_protocol_only(protocol, n, r):
    nodes = _line(n, protocol)
    try:
        init = nodes[0].model.state_dict()["l1.weight"].clone()
        nodes[0].set_start_learning(rounds=r, epochs=0)
        wait_4_results(nodes, 60)
        check_equal_models(nodes, atol=0)
        # epochs=0: everybody ends with exactly the initiator's initial weights
        assert torch.equal(nodes[1].state.learner.get_parameters()["l1.weight"], init)
    finally:
        _stop(nodes)


def test_training_federation_improves_and_agrees():
    nodes = _line(3, InMemoryCommunicationProtocol)
    try:
        logger.reset_metrics()
        nodes[0].set_start_learning(rounds=2, epochs=1)
        wait_4_results(nodes, 120)
        check_equal_models(nodes, atol=1e-6)
        logs = logger.get_global_logs()["experiment"]
        for n in nodes:
            losses = dict(logs[n.addr]["test_loss"])
            assert losses[2] < losses[0]
        # every node's store holds everyone's metrics (they are broadcast)
        assert set(logs) == {n.addr for n in nodes}
        assert logger.get_local_logs()["experiment"]  # per-step train_loss
    finally:
        _stop(nodes)


def test_non_trainers_receive_the_aggregate():
    Settings.TRAIN_SET_SIZE = 2
    nodes = _line(5, InMemoryCommunicationProtocol)
    try:
        nodes[0].set_start_learning(rounds=2, epochs=1)
        wait_4_results(nodes, 120)
        check_equal_models(nodes, atol=1e-6)
        hist = nodes[0].learning_workflow.history
        assert hist[0] == "StartLearningStage" and hist.count("VoteTrainSetStage") == 2  # per-round vote
    finally:
        _stop(nodes)


def test_vote_once_compat_mode():
    Settings.VOTE_EVERY_ROUND = False
    Settings.TRAIN_SET_SIZE = 2
    nodes = _line(3, InMemoryCommunicationProtocol)
    try:
        nodes[0].set_start_learning(rounds=2, epochs=0)
        wait_4_results(nodes, 60)
        assert nodes[0].learning_workflow.history.count("VoteTrainSetStage") == 1
        check_equal_models(nodes, atol=0)
    finally:
        Settings.VOTE_EVERY_ROUND = True
        _stop(nodes)


def test_second_experiment_on_same_nodes():
    """Q2/Q19: a node can run a second experiment."""
    nodes = _line(2, InMemoryCommunicationProtocol)
    try:
        for _ in range(2):
            nodes[0].set_start_learning(rounds=1, epochs=0)
            wait_4_results(nodes, 60)
            assert nodes[0].learning_workflow.finished
            wait_convergence(nodes, 1, only_direct=True)
        check_equal_models(nodes, atol=0)
    finally:
        _stop(nodes)


def test_interrupt_train():
    nodes = _line(2, InMemoryCommunicationProtocol)
    try:
        nodes[0].set_start_learning(rounds=100, epochs=100)
        time.sleep(1.0)
        nodes[0].set_stop_learning()
        wait_4_results(nodes, 30)
        assert all(n.state.round is None for n in nodes)
    finally:
        _stop(nodes)


def test_stop_learning_on_idle_node_is_harmless():
    """Q22."""
    nodes = _line(2, InMemoryCommunicationProtocol)
    try:
        p = nodes[0]._communication_protocol
        p.broadcast(p.build_msg("stop_learning"))
        time.sleep(0.2)
        wait_convergence(nodes, 1, only_direct=True)  # nobody got evicted
    finally:
        _stop(nodes)


@pytest.mark.parametrize("protocol", PROTOCOLS)
@pytest.mark.parametrize("n", [2, 4])
def test_node_down_on_learning(protocol, n):
    nodes = _line(n, protocol)
    try:
        nodes[0].set_start_learning(rounds=2, epochs=0)
        time.sleep(0.3)
        nodes[-1].stop()
        wait_4_results(nodes[:-1], 90)
    finally:
        _stop(nodes[:-1])


@pytest.mark.parametrize("protocol", PROTOCOLS)
def test_wrong_model_stops_receiver(protocol):
    n1 = Node(MLP(), _dm(), protocol=protocol, device="cpu")
    n2 = Node(CNN(), _dm(), protocol=protocol, device="cpu")
    n1.start()
    n2.start()
    try:
        n1.connect(n2.addr)
        time.sleep(0.1)
        n1.set_start_learning(rounds=2, epochs=0)
        wait_4_results([n1, n2], 60)
        assert not n2.is_running()  # the node that got the alien model stopped itself
    finally:
        _stop([n1, n2])


def test_large_model_over_grpc():
    """Q9: the 26 MB CNN does not fit the reference's 4 MiB gRPC default."""
    nodes = _line(2, GrpcCommunicationProtocol, model=CNN)
    try:
        nodes[0].set_start_learning(rounds=1, epochs=0)
        wait_4_results(nodes, 60)
        check_equal_models(nodes, atol=0)
    finally:
        _stop(nodes)


def test_api_guards():
    n = Node(MLP(), _dm(), protocol=InMemoryCommunicationProtocol)
    with pytest.raises(NodeRunningException):
        n.connect("x")
    n.start()
    with pytest.raises(NodeRunningException):
        n.start()
    with pytest.raises(ZeroRoundsException):
        n.set_start_learning(rounds=0)
    n.stop()


def test_tally_votes_is_deterministic():
    ballots = {"a": {"n1": 10, "n2": 5}, "b": {"n2": 5, "n3": 10}, "c": {"n1": 1}}
    # n1=11, n2=10, n3=10 -> ties broken by name descending
    assert tally_votes(ballots, 2) == ["n1", "n3"]
    assert tally_votes(ballots, 5) == ["n1", "n3", "n2"]


def test_aggregator_admission_rules():
    agg = FedAvg("me")
    m = lambda v: {"w": torch.tensor(float(v))}
    assert agg.add_model(m(0), ["a"], 1) == []  # not armed
    agg.set_nodes_to_aggregate(["a", "b", "c"])
    assert agg.add_model(m(0), [], 1) == []  # Q3
    assert agg.add_model(m(1), ["a"], 1) == ["a"]
    assert agg.add_model(m(9), ["a", "b"], 2) == []  # overlaps
    assert agg.add_model(m(9), ["z"], 1) == []  # outsider
    assert sorted(agg.add_model(m(3), ["b", "c"], 3)) == ["a", "b", "c"]
    out = agg.wait_and_get_aggregation(timeout=1)
    assert float(out["w"]) == pytest.approx((1 * 1 + 3 * 3) / 4)
    assert agg.add_model(m(0), ["a"], 1) == []  # finished
    # partial aggregation excludes what the peer already has
    model, contributors, weight = agg.get_partial_aggregation(["a"])
    assert contributors == ["b", "c"] and weight == 3 and float(model["w"]) == 3.0
    assert agg.get_partial_aggregation(["a", "b"]) == (None, None, None)
    agg.clear()
    # full aggregate replaces partials; waiting mode only takes the full set
    agg.set_nodes_to_aggregate(["a", "b"])
    agg.add_model(m(1), ["a"], 1)
    assert sorted(agg.add_model(m(5), ["b", "a"], 7)) == ["a", "b"]
    assert float(agg.wait_and_get_aggregation(timeout=1)["w"]) == 5.0
    agg.clear()
    agg.set_waiting_aggregated_model(["a", "b"])
    assert agg.add_model(m(1), ["a"], 1) == []
    assert agg.add_model(m(2), ["a", "b"], 1) == ["a", "b"]
    assert float(agg.wait_and_get_aggregation(timeout=1)["w"]) == 2.0
    agg.clear()
    # timeout aggregates what is there
    agg.set_nodes_to_aggregate(["a", "b"])
    agg.add_model(m(4), ["a"], 1)
    assert float(agg.wait_and_get_aggregation(timeout=0.05)["w"]) == 4.0


def test_own_model_push_policy():
    """Every trainer sends only its own model: no overlapping partial aggregates, same result."""
    Settings.OWN_MODEL_PUSH = "always"
    Settings.TRAIN_SET_SIZE = 4
    nodes = _line(4, InMemoryCommunicationProtocol)
    try:
        nodes[0].set_start_learning(rounds=2, epochs=1)
        wait_4_results(nodes, 120)
        check_equal_models(nodes, atol=1e-6)
    finally:
        Settings.OWN_MODEL_PUSH = "auto"
        _stop(nodes)


def test_trainer_status_point_to_point():
    """``DIRECT_TRAINSET_STATUS``: `models_aggregated` reaches the round's trainers directly (they are
    not all neighbours on a line) instead of being flooded; the experiment must end the same way."""
    Settings.TRAIN_SET_SIZE = 4  # three peers each, at most two of them neighbours on a line: never all direct
    Settings.DIRECT_TRAINSET_STATUS = True
    nodes = _line(6, InMemoryCommunicationProtocol)
    seen = []
    try:
        for i, n in enumerate(nodes):
            eng = n._communication_protocol
            orig = eng.on_message

            def spy(msg, _orig=orig, _i=i, **kw):
                if msg.cmd == "models_aggregated":
                    seen.append((_i, msg.ttl))
                return _orig(msg, **kw)

            eng.on_message = spy
        nodes[0].set_start_learning(rounds=2, epochs=1)
        wait_4_results(nodes, 120)
        check_equal_models(nodes, atol=1e-6)
        assert seen and all(ttl == 1 for _, ttl in seen), seen[:5]  # nothing was flooded
    finally:
        Settings.DIRECT_TRAINSET_STATUS = False
        _stop(nodes)


def test_early_models_aggregated_notice_is_kept_for_the_next_round():
    """A notice for round r+1 that overtakes this node's own round change is installed when the
    node gets there (dropping it would make the sender look empty-handed for the whole round)."""
    from p2pfl_b200.commands.control import ModelsAggregatedCommand
    from p2pfl_b200.node_state import NodeState

    st = NodeState("me")
    st.set_experiment("experiment", 3)
    cmd = ModelsAggregatedCommand(st)
    cmd.execute("peer", 0, "a", "b")
    assert st.models_aggregated == {"peer": ["a", "b"]}
    cmd.execute("peer", 1, "c")  # next round: stashed
    cmd.execute("other", 5, "z")  # far away: ignored
    assert st.models_aggregated == {"peer": ["a", "b"]}
    st.increase_round()
    assert st.round == 1 and st.models_aggregated == {"peer": ["c"]}
    st.increase_round()
    assert st.models_aggregated == {}


def test_prevote_goes_two_rounds_ahead_only_among_extended_nodes():
    """Ballots are cast ``PREVOTE_AHEAD`` rounds ahead when every known node is one of this framework;
    a reference node files ballots for its current and next round only, so with one in the network the
    pre-vote stays one round ahead.  The handler's window covers what can arrive."""
    from types import SimpleNamespace

    from p2pfl_b200.commands.control import VoteTrainSetCommand
    from p2pfl_b200.node_state import NodeState
    from p2pfl_b200.stages.base_node.train_stage import TrainStage

    Settings.PREVOTE, Settings.VOTE_EVERY_ROUND, Settings.PREVOTE_AHEAD = True, True, 2
    for extended, want in (({"b", "c"}, [1, 2]), ({"b"}, [1])):
        st = NodeState("a")
        st.set_experiment("exp", 10)
        sent = []
        proto = SimpleNamespace(
            _extended=extended,
            get_neighbors=lambda only_direct=False: {"b": None, "c": None},
            build_msg=lambda cmd, args=None, round=None: SimpleNamespace(cmd=cmd, args=args, round=round),
            broadcast_async=lambda msg, lazy=False: sent.append(msg.round),
            broadcast=lambda msg: sent.append(msg.round),
        )
        ctx = SimpleNamespace(state=st, communication_protocol=proto)
        TrainStage.prevote(ctx)
        assert sent == want and sorted(st.train_set_votes) == want
        TrainStage.prevote(ctx)  # idempotent
        assert sent == want
        st.round = 8  # rounds 9 is the last one: nothing beyond it is voted for
        sent.clear()
        TrainStage.prevote(ctx)
        assert sent == [9]
    # window of the handler: a node can be PREVOTE_AHEAD rounds ahead and votes that far ahead of itself
    st = NodeState("a")
    st.set_experiment("exp", 20)
    st.round = 3
    cmd = VoteTrainSetCommand(st)
    for r in (2, 3, 7, 8, 9):
        cmd.execute("b", r, "a", "5")
    assert sorted(st.train_set_votes) == [3, 7, 8]


def test_would_accept_agrees_with_add_model():
    """``Aggregator.would_accept`` (the check that lets a receiver skip decoding / pulling a model it has
    no use for) must answer exactly what ``add_model`` would then do, for any sequence of offers."""
    import copy

    from hypothesis import given, settings
    from hypothesis import strategies as st_

    from p2pfl_b200.learning.aggregators.fedavg import FedAvg

    names = ["a", "b", "c", "d", "e"]
    offers = st_.lists(st_.lists(st_.sampled_from(names + ["x"]), min_size=0, max_size=5, unique=True), min_size=1, max_size=8)

    @settings(max_examples=200, deadline=None)
    @given(train=st_.lists(st_.sampled_from(names), min_size=0, max_size=5, unique=True), waiting=st_.booleans(), seq=offers)
    def check(train, waiting, seq):
        agg = FedAvg(node_name="t")
        if waiting:
            agg.set_waiting_aggregated_model(train)
        else:
            agg.set_nodes_to_aggregate(train)
        for contributors in seq:
            predicted = agg.would_accept(list(contributors))
            before = copy.copy(agg.get_aggregated_models())
            held = agg.add_model({"w": torch.zeros(1)}, list(contributors), 1)
            assert predicted == bool(held), (train, waiting, contributors, before, held)

    check()
