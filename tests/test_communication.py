"""Integration tests of the control plane with real nodes on loopback
(mirrors reference test/communication_test.py) - run over both transports."""

import time

import pytest

from p2pfl_b200.communication.grpc import GrpcCommunicationProtocol
from p2pfl_b200.communication.grpc.address import AddressParser
from p2pfl_b200.communication.memory import InMemoryCommunicationProtocol
from p2pfl_b200.communication.sockets import SocketCommunicationProtocol
from p2pfl_b200.node import Node
from p2pfl_b200.settings import Settings
from p2pfl_b200.utils import full_connection, wait_convergence

PROTOCOLS = [InMemoryCommunicationProtocol, GrpcCommunicationProtocol, SocketCommunicationProtocol]


def _nodes(n, protocol):
    nodes = [Node(None, None, protocol=protocol) for _ in range(n)]
    for x in nodes:
        x.start()
    return nodes


def _stop(nodes):
    for x in nodes:
        x.stop()


@pytest.mark.parametrize("protocol", PROTOCOLS)
def test_connect_invalid_node(protocol):
    n = Node(None, None, protocol=protocol)
    n.start()
    n.connect("google.es:80")
    n.connect("holadani.holaenrique")
    assert len(n.get_neighbors()) == 0
    n.stop()


@pytest.mark.parametrize("protocol", PROTOCOLS)
def test_basic_node_pairing(protocol):
    n1, n2 = _nodes(2, protocol)
    try:
        assert n1.connect(n2.addr)
        wait_convergence([n1, n2], 1, only_direct=True)
        n2.disconnect(n1.addr)
        wait_convergence([n1, n2], 0, only_direct=True, wait=Settings.HEARTBEAT_PERIOD * 2 + 1)
    finally:
        _stop([n1, n2])


@pytest.mark.parametrize("protocol", PROTOCOLS)
def test_full_connected(protocol):
    n1, n2, n3, n4 = nodes = _nodes(4, protocol)
    try:
        n1.connect(n2.addr)
        wait_convergence([n1, n2], 1, only_direct=True)
        full_connection(n3, [n1, n2])
        wait_convergence([n1, n2, n3], 2, only_direct=True)
        full_connection(n4, [n1, n2, n3])
        wait_convergence(nodes, 3, only_direct=True)
        n1.stop()
        wait_convergence([n2, n3, n4], 2, only_direct=True)
        n2.stop()
        wait_convergence([n3, n4], 1, only_direct=True)
        n3.stop()
        wait_convergence([n4], 0, only_direct=True)
    finally:
        _stop(nodes)


@pytest.mark.parametrize("protocol", PROTOCOLS)
def test_network_neighbors_star(protocol):
    """Indirect discovery through TTL-flooded heartbeats on a star."""
    nodes = _nodes(6, protocol)
    hub, leaves = nodes[0], nodes[1:]
    try:
        for leaf in leaves:
            leaf.connect(hub.addr)
        wait_convergence(nodes, 5, only_direct=False, wait=6)
        assert len(hub.get_neighbors(only_direct=True)) == 5
        assert all(len(l.get_neighbors(only_direct=True)) == 1 for l in leaves)
        leaves[-1].stop()
        wait_convergence(nodes[:-1], 4, only_direct=False, wait=8)
    finally:
        _stop(nodes)


@pytest.mark.parametrize("protocol", PROTOCOLS)
def test_bad_msg_evicts(protocol):
    n1, n2 = _nodes(2, protocol)
    try:
        n1.connect(n2.addr)
        wait_convergence([n1, n2], 1, only_direct=True)
        proto = n1._communication_protocol
        deadline = time.time() + 5  # commands outside the reference's set only go to peers known to speak them
        while n2.addr not in proto._extended and time.time() < deadline:
            time.sleep(0.01)
        proto.broadcast(proto.build_msg("BAD_MSG"))
        wait_convergence([n1, n2], 0, only_direct=True, wait=4)
    finally:
        _stop([n1, n2])


@pytest.mark.parametrize("protocol", PROTOCOLS)
def test_node_abrupt_down(protocol):
    """Fault injection: kill only the heartbeater of one node, only the server
    of another; the survivors must evict them."""
    n1, n2, n3 = nodes = _nodes(3, protocol)
    try:
        n1.connect(n2.addr)
        n3.connect(n1.addr)
        n3.connect(n2.addr)
        wait_convergence(nodes, 2, only_direct=True)
        n1._communication_protocol._heartbeater.stop()
        wait_convergence([n2, n3], 1, only_direct=False, wait=10)
        n2._communication_protocol._server.stop()
        wait_convergence([n3], 0, only_direct=False, wait=10)
    finally:
        _stop(nodes)


def test_ttl_bounds_reach_and_relay_copies_message():
    """A message floods exactly TTL hops down a line, and the relay works on a
    copy (the reference decrements the shared dict, Q6)."""
    Settings.TTL = 3
    nodes = _nodes(6, InMemoryCommunicationProtocol)
    seen = []
    try:
        from p2pfl_b200.commands.command import Command

        class Probe(Command):
            def __init__(self, who):
                self.who = who

            @staticmethod
            def get_name():
                return "probe"

            def execute(self, source, round, *args, **kw):
                seen.append(self.who)

        for i, n in enumerate(nodes):
            n._communication_protocol.add_command(Probe(i))
        for i in range(5):
            nodes[i + 1].connect(nodes[i].addr)
        wait_convergence(nodes[1:5], 2, only_direct=True)
        time.sleep(0.5)  # greeting beats: every node learns its neighbours speak the extended commands
        proto = nodes[0]._communication_protocol
        msg = proto.build_msg("probe")
        proto.broadcast(msg)
        time.sleep(0.5)
        assert sorted(seen) == [1, 2, 3]
        assert msg.ttl == 3  # original untouched
    finally:
        _stop(nodes)


def test_duplicate_filter_ring():
    Settings.AMOUNT_LAST_MESSAGES_SAVED = 3
    n = Node(None, None, protocol=InMemoryCommunicationProtocol)
    g = n._communication_protocol._gossiper
    assert g.check_and_set_processed(1) and not g.check_and_set_processed(1)
    for h in (2, 3, 4):
        assert g.check_and_set_processed(h)
    assert g.check_and_set_processed(1)  # fell out of the ring


def test_address_parser():
    assert AddressParser("127.0.0.1:5000").get_parsed_address() == "127.0.0.1:5000"
    assert AddressParser("[::1]:5000").get_parsed_address() == "[::1]:5000"
    assert AddressParser("unix:///tmp/x.sock").get_parsed_address() == "unix:///tmp/x.sock"
    bare = AddressParser("127.0.0.1")
    assert bare.port and bare.get_parsed_address() == f"127.0.0.1:{bare.port}"
    with pytest.raises(ValueError):
        AddressParser("host:notaport")


def test_proto_wire_bytes():
    """Field numbers/types match node.proto (hand-checked protobuf encoding)."""
    from p2pfl_b200.communication.grpc import proto

    pb = proto.PbMessage(source="a", ttl=2, hash=5, cmd="c", args=["x"])
    pb.round = 7
    assert pb.SerializeToString() == b"\x0a\x01a\x10\x02\x18\x05\x22\x01c\x2a\x01x\x30\x07"
    w = proto.PbWeights(source="s", round=1, weights=b"\x00\x01", contributors=["p", "q"], weight=3, cmd="add_model")
    assert w.SerializeToString() == b"\x0a\x01s\x10\x01\x1a\x02\x00\x01\x22\x01p\x22\x01q\x28\x03\x32\x09add_model"
    assert not proto.PbResponseMessage().HasField("error")


def test_unix_socket_transport(tmp_path):
    a = Node(None, None, address=f"unix://{tmp_path}/a.sock", protocol=GrpcCommunicationProtocol)
    b = Node(None, None, address=f"unix://{tmp_path}/b.sock", protocol=GrpcCommunicationProtocol)
    a.start()
    b.start()
    try:
        assert a.connect(b.addr)
        wait_convergence([a, b], 1, only_direct=True)
    finally:
        _stop([a, b])


def test_relay_pruning_in_a_clique_and_reach_on_a_line():
    """In a full mesh a broadcast costs N-1 deliveries (peers advertise their direct
    neighbours in their beats, so nobody relays); on a line it still floods end to end."""
    from p2pfl_b200.commands.command import Command

    class Probe(Command):
        hits = []

        def __init__(self, who):
            self.who = who

        @staticmethod
        def get_name():
            return "probe2"

        def execute(self, source, round, *args, **kw):
            Probe.hits.append(self.who)

    for topology in ("clique", "line"):
        nodes = _nodes(4, InMemoryCommunicationProtocol)
        deliveries = []
        try:
            for i, n in enumerate(nodes):
                n._communication_protocol.add_command(Probe(i))
                eng = n._communication_protocol
                orig = eng.on_message

                def counted(msg, _orig=orig, _i=i, **kw):
                    if msg.cmd == "probe2":
                        deliveries.append(_i)
                    return _orig(msg, **kw)

                eng.on_message = counted
            if topology == "clique":
                for i in range(4):
                    for j in range(i):
                        nodes[i].connect(nodes[j].addr)
                wait_convergence(nodes, 3, only_direct=True)
            else:
                for i in range(3):
                    nodes[i + 1].connect(nodes[i].addr)
                wait_convergence(nodes, 3, only_direct=False)
            time.sleep(Settings.HEARTBEAT_PERIOD * 2.5)  # let the beats advertise neighbour lists
            Probe.hits = []
            p = nodes[0]._communication_protocol
            p.broadcast(p.build_msg("probe2"))
            time.sleep(0.5)
            assert sorted(Probe.hits) == [1, 2, 3], (topology, Probe.hits)
            if topology == "clique":
                assert len(deliveries) == 3, deliveries  # no redundant relays
        finally:
            _stop(nodes)


@pytest.mark.parametrize("protocol", [InMemoryCommunicationProtocol, SocketCommunicationProtocol])
def test_direct_neighbours_are_advertised_when_they_change(protocol):
    """A freshly connected mesh must not flood for a heartbeat period: the list of direct neighbours
    goes out the moment it (or the set of peers that understand it) changes, versioned so that an
    overtaken older list cannot replace a newer one."""
    Settings.HEARTBEAT_PERIOD = 30.0  # the periodic hint never comes within this test
    Settings.HEARTBEAT_TIMEOUT = 120.0
    nodes = _nodes(4, protocol)
    try:
        for i in range(4):
            for j in range(i):
                nodes[i].connect(nodes[j].addr)
        wait_convergence(nodes, 3, only_direct=True)
        addrs = {n.addr for n in nodes}

        def settled():
            return all(
                n._communication_protocol._direct_of.get(o.addr) == addrs - {o.addr}
                for n in nodes for o in nodes if o is not n
            )  # fmt: skip

        deadline = time.time() + 3
        while not settled() and time.time() < deadline:
            time.sleep(0.01)
        assert settled(), [n._communication_protocol._direct_of for n in nodes]
        # an older hint arriving late is ignored
        eng = nodes[0]._communication_protocol
        v = eng._direct_of_version[nodes[1].addr]
        eng.note_direct_neighbors_of(nodes[1].addr, [], v - 1)
        assert eng._direct_of[nodes[1].addr] == addrs - {nodes[1].addr}
        # a peer that leaves is advertised as gone at once, and its own hints are forgotten
        gone = nodes[3].addr
        nodes[3].stop()
        deadline = time.time() + 3
        while time.time() < deadline and any(
            gone in n._communication_protocol._direct_of.get(o.addr, ()) for n in nodes[:3] for o in nodes[:3] if o is not n
        ):  # fmt: skip
            time.sleep(0.01)
        for n in nodes[:3]:
            d = n._communication_protocol._direct_of
            assert gone not in d and all(gone not in s for s in d.values()), d
    finally:
        _stop(nodes)


@pytest.mark.parametrize("protocol", PROTOCOLS)
def test_bookkeeping_broadcasts_are_bundled(protocol):
    """A lazily queued broadcast waits for the next ordinary one; together they reach an extended
    neighbour as ONE message, a neighbour that has not shown the capability mark one by one."""
    # only the greeting beats of the connects: a periodic beat of node 2 would re-mark it as extended
    # (and its hint would carry the lazily queued messages away) in the middle of the scenario
    Settings.HEARTBEAT_PERIOD = 30.0
    Settings.HEARTBEAT_TIMEOUT = 120.0
    nodes = _nodes(3, protocol)
    got = {1: [], 2: []}
    try:
        from p2pfl_b200.commands.command import Command

        class Metrics(Command):
            def __init__(self, who):
                self.who = who

            @staticmethod
            def get_name():
                return "metrics"

            def execute(self, source, round, *args, **kw):
                got[self.who].append(("metrics", round, args))

        class Ready(Metrics):
            @staticmethod
            def get_name():
                return "models_ready"

            def execute(self, source, round, *args, **kw):
                got[self.who].append(("models_ready", round, args))

        for i in (1, 2):
            nodes[i]._communication_protocol.add_command([Metrics(i), Ready(i)])
        nodes[0].connect(nodes[1].addr)
        nodes[0].connect(nodes[2].addr)
        wait_convergence(nodes[:1], 2, only_direct=True)
        proto = nodes[0]._communication_protocol
        deadline = time.time() + 5
        while len(proto._extended) < 2 and time.time() < deadline:
            time.sleep(0.01)
        proto._extended.discard(nodes[2].addr)  # node 2 plays a reference node
        # the neighbour lists advertised when the two peers showed up leave first (possibly bundled)
        deadline = time.time() + 5
        while proto._gossiper._pending and time.time() < deadline:
            time.sleep(0.01)
        time.sleep(0.2)
        sent = []
        real = proto._transport.deliver_many

        def spy(targets, env):
            sent.extend((t[0], env.cmd) for t in targets)
            return real(targets, env)

        proto._transport.deliver_many = spy
        proto.broadcast_async(proto.build_msg("models_ready", [], round=4), lazy=True)
        proto.broadcast_async(proto.build_msg("metrics", ["test_loss", "0.5"], round=5), lazy=True)
        time.sleep(Settings.BUNDLE_LINGER / 4)
        assert not got[1] and not got[2]  # still waiting for company
        proto.broadcast_async(proto.build_msg("metrics", ["test_metric", "0.9"], round=5))
        deadline = time.time() + 5
        while (len(got[1]) < 3 or len(got[2]) < 3) and time.time() < deadline:
            time.sleep(0.01)
        want = [("models_ready", 4, ()), ("metrics", 5, ("test_loss", "0.5")), ("metrics", 5, ("test_metric", "0.9"))]
        assert got[1] == want and got[2] == want
        mine = [c for n, c in sent if c != "beat" and c != "neighbors_hint"]
        assert sorted(c for n, c in sent if n == nodes[1].addr and c in mine) == ["bundle"]
        assert sorted(c for n, c in sent if n == nodes[2].addr and c in mine) == ["metrics", "metrics", "models_ready"]
        # a lone lazy broadcast leaves by itself after the linger
        proto.broadcast_async(proto.build_msg("models_ready", [], round=6), lazy=True)
        deadline = time.time() + 5
        while len(got[1]) < 4 and time.time() < deadline:
            time.sleep(0.01)
        assert got[1][-1] == ("models_ready", 6, ())
    finally:
        _stop(nodes)


def test_bundled_relays_keep_their_originator():
    """A relayer forwards several messages to one neighbour as one bundle; the receiver must see
    each with its original source, TTL and id (and drop the ones it already has)."""
    nodes = _nodes(3, InMemoryCommunicationProtocol)
    got = []
    try:
        from p2pfl_b200.commands.command import Command

        class Probe(Command):
            @staticmethod
            def get_name():
                return "probe"

            def execute(self, source, round, *args, **kw):
                got.append((source, round, args))

        nodes[2]._communication_protocol.add_command(Probe())
        nodes[1].connect(nodes[2].addr)
        wait_convergence(nodes[1:], 1, only_direct=True)
        relayer = nodes[1]._communication_protocol
        deadline = time.time() + 5
        while nodes[2].addr not in relayer._extended and time.time() < deadline:
            time.sleep(0.01)
        origin = nodes[0]._communication_protocol
        msgs = [origin.build_msg("probe", [str(i)], round=i) for i in range(3)]
        relayer._send_bundle_to(nodes[2].addr, [m.relay_copy() for m in msgs])
        relayer._send_bundle_to(nodes[2].addr, [m.relay_copy() for m in msgs[:2]])  # duplicates: dropped
        assert got == [(nodes[0].addr, i, (str(i),)) for i in range(3)]
    finally:
        _stop(nodes)


def test_rendered_proto_is_current():
    """The ``.proto`` text shipped for other languages is rendered from the descriptors in use."""
    import importlib.util
    import os

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("render_proto", os.path.join(root, "tools", "render_proto.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    with open(os.path.join(root, "p2pfl_b200", "communication", "grpc", "proto_schema", "wire_schema.proto")) as fh:
        assert fh.read() == mod.render()


def test_socket_transport_reaches_nodes_of_its_own_process_directly():
    """Several simulated nodes per process (config 4): a neighbour that lives in this process is reached
    by a call into its engine, not over the loopback socket; a node that goes down is noticed the same
    way, and a learning experiment over such links ends with equal models."""
    from p2pfl_b200.communication.sockets.socket_communication_protocol import _LocalLink
    from p2pfl_b200.learning.data import MnistFederatedDM
    from p2pfl_b200.models import MLP
    from p2pfl_b200.utils import check_equal_models, wait_4_results

    Settings.SOCKET_LOCAL_SHORTCUT = True
    nodes = [Node(MLP(seed=i), MnistFederatedDM(i, 3, n_train=96 * 3, n_test=32 * 3), protocol=SocketCommunicationProtocol, device="cpu") for i in range(3)]
    for n in nodes:
        n.start()
    try:
        nodes[1].connect(nodes[0].addr)
        nodes[2].connect(nodes[1].addr)  # a line: 0 - 1 - 2
        wait_convergence(nodes, 2, only_direct=False)
        for n in nodes:
            links = [e.link for e in n._communication_protocol.get_neighbors(only_direct=True).values()]
            assert links and all(isinstance(l, _LocalLink) for l in links), links
        nodes[0].set_start_learning(rounds=2, epochs=1)
        wait_4_results(nodes)
        check_equal_models(nodes)
        # a node that stops is evicted by the next send, as over the wire
        nodes[2]._communication_protocol._server.stop()
        p1 = nodes[1]._communication_protocol
        deadline = time.time() + 10
        while nodes[2].addr in p1.get_neighbors(only_direct=True) and time.time() < deadline:
            p1.broadcast(p1.build_msg("beat", [str(time.time())]))
            time.sleep(0.05)
        assert nodes[2].addr not in p1.get_neighbors(only_direct=True)
    finally:
        _stop(nodes)
