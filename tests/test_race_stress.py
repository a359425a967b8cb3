"""Race / lock-order stress of the control plane (the reference has no such tooling).

Every lock of the framework is created through ``utils.lockcheck``; with checking on, each
acquisition records which locks the thread already holds.  A cycle in the resulting order graph is a
deadlock two threads can realise.  The tests churn real nodes from several threads - connects,
disconnects, floods of control messages, a node dying mid-experiment - and then assert that no cycle
was seen, no thread died with an exception and the survivors finished learning."""

import random
import threading
import time

import pytest

from p2pfl_b200.utils import lockcheck


@pytest.fixture
def checked_locks():
    lockcheck.enable(True)
    lockcheck.reset()
    errors = []
    old_hook = threading.excepthook
    threading.excepthook = lambda a: errors.append((a.thread.name if a.thread else "?", repr(a.exc_value)))
    yield errors
    threading.excepthook = old_hook
    lockcheck.enable(False)


def _protocols():
    from p2pfl_b200.communication.memory import InMemoryCommunicationProtocol
    from p2pfl_b200.communication.sockets import SocketCommunicationProtocol

    return [InMemoryCommunicationProtocol, SocketCommunicationProtocol]


def test_lockcheck_detects_an_inversion():
    lockcheck.enable(True)
    lockcheck.reset()
    try:
        a, b = lockcheck.make_lock("test.a"), lockcheck.make_lock("test.b")
        with a:
            with b:
                pass
        assert lockcheck.violations() == []
        with b:
            with a:  # opposite order: never deadlocks here (one thread), but two threads could
                pass
        v = lockcheck.violations()
        assert len(v) == 1 and set(v[0]) == {"test.a", "test.b"}
    finally:
        lockcheck.enable(False)
        lockcheck.reset()


@pytest.mark.parametrize("which", [0, 1])
def test_connection_churn_and_message_floods(checked_locks, which):
    from p2pfl_b200.commands.command import Command
    from p2pfl_b200.node import Node

    proto = _protocols()[which]
    hits = []

    class Probe(Command):
        @staticmethod
        def get_name():
            return "probe"

        def execute(self, source, round, *args, **kw):
            hits.append(source)

    nodes = [Node(None, None, protocol=proto) for _ in range(8)]
    for n in nodes:
        n._communication_protocol.add_command(Probe())
        n.start()
    stop = threading.Event()
    rng = random.Random(7)

    def churn(seed):
        r = random.Random(seed)
        while not stop.is_set():
            a, b = r.sample(nodes, 2)
            try:
                if r.random() < 0.6:
                    a.connect(b.addr)
                else:
                    a.disconnect(b.addr)
            except Exception:
                pass
            time.sleep(r.random() * 0.004)

    def flood(seed):
        r = random.Random(seed)
        while not stop.is_set():
            n = r.choice(nodes)
            p = n._communication_protocol
            try:
                p.broadcast(p.build_msg("probe", [str(r.random())]))
            except Exception:
                pass
            time.sleep(r.random() * 0.002)

    threads = [threading.Thread(target=churn, args=(s,), name=f"churn-{s}") for s in range(3)]
    threads += [threading.Thread(target=flood, args=(10 + s,), name=f"flood-{s}") for s in range(3)]
    try:
        for t in threads:
            t.start()
        time.sleep(3.0)
    finally:
        stop.set()
        for t in threads:
            t.join(10)
        for n in nodes:
            n.stop()
    assert not any(t.is_alive() for t in threads), "a stress thread is stuck (deadlock?)"
    assert checked_locks == [], checked_locks
    assert lockcheck.violations() == [], lockcheck.violations()
    assert hits, "no probe message was ever delivered"
    del rng


def test_learning_with_a_node_dying_midway(checked_locks):
    """An experiment over the in-memory transport while one trainer is stopped abruptly: the others
    evict it and finish; no lock-order cycle, no crashed thread."""
    from p2pfl_b200.communication.memory import InMemoryCommunicationProtocol
    from p2pfl_b200.learning.data import MnistFederatedDM
    from p2pfl_b200.learning.torch_learner import TorchLearner
    from p2pfl_b200.models import MLP
    from p2pfl_b200.node import Node
    from p2pfl_b200.settings import Settings
    from p2pfl_b200.utils import wait_convergence

    Settings.TRAIN_SET_SIZE = 4
    n = 5
    nodes = [
        Node(MLP(seed=i), MnistFederatedDM(i, n, n_train=64 * n, n_test=32 * n), protocol=InMemoryCommunicationProtocol,
             learner=TorchLearner, device="cpu")
        for i in range(n)
    ]  # fmt: skip
    for x in nodes:
        x.start()
    try:
        for i in range(1, n):
            for j in range(i):
                nodes[i].connect(nodes[j].addr)
        wait_convergence(nodes, n - 1, only_direct=True)
        nodes[0].set_start_learning(rounds=3, epochs=1)
        deadline = time.time() + 30
        while nodes[0].state.round in (None, 0) and time.time() < deadline:
            time.sleep(0.01)
        nodes[-1].stop()  # dies after round 0
        survivors = nodes[:-1]
        for x in survivors:
            assert x.wait_learning_finished(120), f"{x.addr} did not finish after its peer died"
    finally:
        for x in nodes:
            x.stop()
    assert checked_locks == [], checked_locks
    assert lockcheck.violations() == [], lockcheck.violations()
