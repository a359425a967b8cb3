"""The decision "this round is exchanged by the fused kernel" must come out the same on every
trainer (a trainer that skips it would leave the others spinning on flags nobody writes).  It is
made from the train set and the members' arena announcements; these tests drive it on the CPU with
stand-in arenas (no device is touched before a plan is launched)."""

import os
import socket
import threading
import time
from types import SimpleNamespace

from p2pfl_b200.communication.nvlink import NvlinkMemoryProtocol
from p2pfl_b200.node_state import NodeState
from p2pfl_b200.parallel.arena import ArenaGeometry

GEO = ArenaGeometry(235168, 8, 16384, True)
HOST, PID = socket.gethostname(), os.getpid()


def _info(device=0, pid=PID, fused=1, geo=GEO, layout="abc", host=HOST):
    return [geo.key(), layout, str(device), str(pid), host, "00" * 64, "100", str(fused)]


def _proto(addr, train_set, device=0):
    p = NvlinkMemoryProtocol(addr)
    st = NodeState(p.addr)
    st.round, st.train_set = 0, list(train_set)
    learner = SimpleNamespace(
        arena=SimpleNamespace(geo=GEO, device=device), layout=SimpleNamespace(hash="abc"), get_num_samples=lambda: (100, 10),
    )  # fmt: skip
    node = SimpleNamespace(state=st)
    p.bind_node(node)
    p._learner = lambda: learner  # type: ignore[method-assign]
    p._my_peerinfo = __import__("p2pfl_b200.communication.nvlink.nvlink_protocol", fromlist=["PeerInfo"]).PeerInfo(_info(device))
    p._map_peer = lambda a: SimpleNamespace(addr=a)  # type: ignore[method-assign]
    p.get_neighbors = lambda only_direct=False: {a: None for a in train_set}  # type: ignore[method-assign]
    sent = []
    p.broadcast_async = lambda msg, lazy=False: sent.append(msg)  # type: ignore[method-assign]
    return p, st, learner, sent


def test_every_member_willing_gives_a_plan_with_the_local_group():
    ts = ["a", "b", "c"]
    p, st, learner, _ = _proto("a", ts)
    p._on_arena_info("b", _info(device=0))          # same process, same GPU: one cooperative launch with me
    p._on_arena_info("c", _info(device=1))          # another GPU
    plan = p.plan_fused_exchange(st, learner, wait=0.5)
    assert plan is not None and plan.local == ["a", "b"] and not plan.can_prelaunch
    assert plan.train_set == ts and set(plan.peers) == {"b", "c"}


def test_an_unwilling_or_mismatching_member_means_no_plan_for_anybody():
    ts = ["a", "b", "c"]
    for bad in (_info(fused=0), _info(layout="other"), _info(geo=ArenaGeometry(64, 8, 16384, True)), _info(host="elsewhere")):
        answers = []
        for me in ts:  # every trainer evaluates the same announcements -> the same answer
            p, st, learner, _ = _proto(me, ts)
            for other in ts:
                if other != me:
                    p._on_arena_info(other, bad if other == "c" or me == "c" else _info(device=ts.index(other)))
            if me == "c":
                p._my_peerinfo.fused = False if bad[7] == "0" else p._my_peerinfo.fused
            answers.append(p.plan_fused_exchange(st, learner, wait=0.2) is not None)
        assert answers[0] is False and answers[1] is False, (bad, answers)


def test_a_missing_announcement_is_waited_for_not_guessed():
    ts = ["a", "b"]
    p, st, learner, _ = _proto("a", ts)
    threading.Timer(0.3, lambda: p._on_arena_info("b", _info(device=1))).start()
    t0 = time.time()
    plan = p.plan_fused_exchange(st, learner, wait=5.0)
    assert plan is not None and 0.25 < time.time() - t0 < 3.0 and plan.can_prelaunch
    # and without the announcement: no plan once the wait is over
    q, st2, learner2, _ = _proto("a", ts)
    assert q.plan_fused_exchange(st2, learner2, wait=0.2) is None


def test_trainers_of_another_process_on_my_gpu_rule_the_fused_exchange_out():
    ts = ["a", "b"]
    p, st, learner, _ = _proto("a", ts)
    p._on_arena_info("b", _info(device=0, pid=PID + 1))
    assert p.plan_fused_exchange(st, learner, wait=0.2) is None


def test_a_local_mapping_failure_is_announced_and_a_cancellation_is_honoured():
    ts = ["a", "b"]
    p, st, learner, sent = _proto("a", ts)
    p._on_arena_info("b", _info(device=1))
    p._map_peer = lambda a: None  # type: ignore[method-assign]
    assert p.plan_fused_exchange(st, learner, wait=0.2) is None
    assert [m.cmd for m in sent] == ["fused_cancel"] and sent[0].round == 0 and p.fused_cancelled(0)
    # the peer that receives the cancellation skips the exchange of that round, not of the next
    q, st2, learner2, _ = _proto("b", ts)
    q._on_arena_info("a", _info(device=0))
    q._commands["fused_cancel"].execute("a", 0)
    assert q.plan_fused_exchange(st2, learner2, wait=0.2) is None
    st2.round = 1
    assert q.plan_fused_exchange(st2, learner2, wait=0.2) is not None
    # a cancellation that arrives after a completed exchange re-opens the diffusion to that trainer
    st2.fused_complete_round = 1
    q._commands["fused_cancel"].execute("a", 1)
    assert st2.fused_complete_round is None


def test_arena_announcement_reaches_a_peer_before_the_first_device_reference():
    """An ``arena_info`` broadcast skips peers not yet known to speak the extended commands (found on
    two GPUs: rank 1's greeting beat was 200 ms late, rank 0 announced meanwhile, and the first device
    reference rank 1 received named an arena it had never heard of).  Whatever the timing: the sender
    delivers its announcement ahead of the first reference on the same link, and re-announces to a peer
    that shows the mark later."""
    import torch

    from p2pfl_b200.commands.command import Command
    from p2pfl_b200.communication.nvlink import NvlinkSocketProtocol
    from p2pfl_b200.learning.params import FlatParams
    from p2pfl_b200.parallel.payload import DevicePayload, DeviceRef

    def learner_for(tag):
        params = torch.zeros(GEO.n)
        arena = SimpleNamespace(geo=GEO, device=0, params=params, outbox=torch.zeros(GEO.n), ipc_handle=lambda: bytes([tag]) * 64)
        return SimpleNamespace(arena=arena, layout=SimpleNamespace(hash="abc", total=GEO.n), get_num_samples=lambda: (100, 10), flat=params)

    a, b = NvlinkSocketProtocol("127.0.0.1"), NvlinkSocketProtocol("127.0.0.1")
    arrivals = []

    class Init(Command):
        @staticmethod
        def get_name():
            return "init_model"

        def execute(self, source, round, *args, weights=None, **kw):
            arrivals.append(("init_model", DeviceRef.is_ref(weights), source in b._peers))

    for p in (a, b):
        p.bind_node(SimpleNamespace(state=NodeState(p.addr)))
    b.add_command(Init())
    real_info = b._on_arena_info
    b._on_arena_info = lambda src, args: (arrivals.append(("arena_info", src)), real_info(src, args))[1]  # type: ignore[method-assign]
    a.start()
    b.start()
    try:
        assert b.connect(a.addr)
        deadline = time.time() + 5
        while not (b.addr in a._extended and a.addr in b._extended) and time.time() < deadline:
            time.sleep(0.01)
        la, lb = learner_for(1), learner_for(2)
        a._extended.discard(b.addr)  # "b's greeting beat has not arrived yet"
        a.publish_arena(la)
        time.sleep(0.1)
        assert a.addr not in b._peers and not arrivals  # the broadcast skipped b
        a._extended.add(b.addr)
        b.publish_arena(lb)  # a learns that b has an arena -> it will send references
        assert b.addr in a._peers
        payload = DevicePayload(FlatParams(la.flat, la.layout, {}), la)
        a.send(b.addr, a.build_weights("init_model", 0, payload))
        assert arrivals == [("arena_info", a.addr), ("init_model", True, True)], arrivals
        # a second reference does not repeat the announcement
        a.send(b.addr, a.build_weights("init_model", 0, payload))
        assert [x[0] for x in arrivals] == ["arena_info", "init_model", "init_model"]
        # a peer that shows the mark after the announcement gets its copy
        arrivals.clear()
        a._extended.discard(b.addr)
        a._note_extended(b.addr)
        deadline = time.time() + 5
        while not arrivals and time.time() < deadline:
            time.sleep(0.01)
        assert arrivals == [("arena_info", a.addr)]
    finally:
        a.stop()
        b.stop()
