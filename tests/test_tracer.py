"""Host timeline tracer (management/tracer.py): spans of stages, message handlers and sends of a real
two-node experiment end up in a Chrome trace-event file; switched off it records nothing."""

import json

from p2pfl_b200.communication.memory import InMemoryCommunicationProtocol
from p2pfl_b200.learning.data import MnistFederatedDM
from p2pfl_b200.management import tracer
from p2pfl_b200.models import MLP
from p2pfl_b200.node import Node
from p2pfl_b200.utils import wait_4_results, wait_convergence


def _experiment():
    nodes = [Node(MLP(seed=i), MnistFederatedDM(i, 2, n_train=256, n_test=64), protocol=InMemoryCommunicationProtocol, device="cpu") for i in range(2)]
    try:
        for n in nodes:
            n.start()
        nodes[1].connect(nodes[0].addr)
        wait_convergence(nodes, 1, only_direct=True)
        nodes[0].set_start_learning(rounds=2, epochs=1)
        wait_4_results(nodes)
        return [n.addr for n in nodes]
    finally:
        for n in nodes:
            n.stop()


def test_trace_of_an_experiment(tmp_path):
    tracer.reset()
    tracer.enable(str(tmp_path / "trace_{pid}.json"))
    try:
        addrs = _experiment()
        path = tracer.dump()
    finally:
        tracer.enable(None)
    assert path is not None and "{pid}" not in path
    doc = json.load(open(path))
    ev = doc["traceEvents"]
    procs = {e["args"]["name"]: e["pid"] for e in ev if e.get("name") == "process_name"}
    assert set(addrs) <= set(procs)
    spans = [e for e in ev if e.get("ph") == "X"]
    assert all(e["dur"] >= 0 and e["ts"] >= 0 for e in spans)
    for addr in addrs:
        mine = [e for e in spans if e["pid"] == procs[addr]]
        stages = {e["name"] for e in mine if e["cat"] == "stage"}
        assert {"StartLearningStage", "VoteTrainSetStage", "TrainStage", "GossipModelStage", "RoundFinishedStage"} <= stages
        assert sum(1 for e in mine if e["cat"] == "stage" and e["name"] == "TrainStage") == 2  # one per round
        assert any(e["cat"] == "message" and e["name"] == "vote_train_set" for e in mine)
        assert any(e["cat"] == "phase" and e["name"] == "fit" and e["dur"] > 0 for e in mine)
        assert any(e["cat"] == "send" for e in mine)
    assert any(e["cat"] == "model" for e in spans)  # the initial weights and the partial aggregates
    assert any(e.get("name") == "thread_name" for e in ev)
    assert "stage:TrainStage" in tracer.summary()
    tracer.reset()


def test_tracer_off_records_nothing():
    tracer.reset()
    assert not tracer.enabled()
    with tracer.span("x", "n"):
        pass
    tracer.instant("y", "n")
    tracer.complete("z", "n", 0.0, 1.0)
    assert tracer.summary() == {} and tracer.dump("/nonexistent/should/not/be/written.json") is None
