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


def test_round_switch_against_inbound_votes_and_notices():
    """Ballots and ``models_aggregated`` notices of faster peers arrive on handler threads while the
    learning thread switches rounds (and prunes the same tables).  Found with
    bench/control_plane_bench.py: one run in ~20 of 8 nodes x 100 rounds died with "dictionary changed
    size during iteration" in ``increase_round`` and left the other seven waiting for it."""
    from p2pfl_b200.commands.control import ModelsAggregatedCommand, VoteTrainSetCommand
    from p2pfl_b200.node_state import NodeState

    st = NodeState("me")
    st.set_experiment("exp", 10**9)
    vote, agg = VoteTrainSetCommand(st), ModelsAggregatedCommand(st)
    stop = threading.Event()
    errors = []

    def handler(i: int) -> None:
        try:
            k = 0
            while not stop.is_set():
                r = st.round or 0
                vote.execute(f"peer{i}", r + 1 + (k % 3), "a", "1", "b", "2")
                agg.execute(f"peer{i}", r + 1, "a", "b")
                k += 1
        except Exception as e:  # pragma: no cover - the failure being tested for
            errors.append(e)

    import sys

    old_interval = sys.getswitchinterval()
    sys.setswitchinterval(1e-6)  # switch threads at (nearly) every opportunity
    threads = [threading.Thread(target=handler, args=(i,), daemon=True) for i in range(4)]
    for t in threads:
        t.start()
    try:
        t_end = time.time() + 0.5
        while time.time() < t_end and not errors:
            st.increase_round()
    except Exception as e:
        errors.append(e)
    finally:
        stop.set()
        for t in threads:
            t.join(5)
        sys.setswitchinterval(old_interval)
    assert not errors, errors
    assert all(r >= st.round for r in st.train_set_votes)  # pruned, and nothing filed behind the switch


def test_round_switch_holds_the_lock_the_handlers_take():
    """The same race, made deterministic: a ballot for a new round is filed by another thread exactly
    while ``increase_round`` walks the vote table.  With the table walked outside the state's lock the
    insertion lands mid-iteration (RuntimeError); with it inside, the handler waits for the switch."""
    from p2pfl_b200.commands.control import VoteTrainSetCommand
    from p2pfl_b200.node_state import NodeState

    st = NodeState("me")
    st.set_experiment("exp", 100)
    st.round = 5
    vote = VoteTrainSetCommand(st)
    filed = []

    class PokingDict(dict):
        def __iter__(self):
            first = True
            for k in dict.__iter__(self):
                yield k
                if first:
                    first = False
                    t = threading.Thread(target=lambda: (vote.execute("peer", st.round + 3, "a", "1"), filed.append(1)), daemon=True)
                    t.start()
                    t.join(0.3)  # old code: the ballot is in by now; new code: the handler is blocked

    st.train_set_votes = PokingDict({4: {"x": {}}, 5: {"x": {}}, 6: {"x": {}}})
    st.increase_round()  # must not raise
    deadline = time.time() + 5
    while not filed and time.time() < deadline:
        time.sleep(0.01)
    assert filed and sorted(st.train_set_votes) == [6, 9]
