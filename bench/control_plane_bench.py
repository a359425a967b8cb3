#!/usr/bin/env python
"""Host-side cost of a federated round, measured WITHOUT a GPU.

Real nodes, stage machine and socket control plane - one process per "GPU", ``--per-proc`` nodes in
each - and an emulated device: ``fit`` sleeps for the device time of the round (the GIL is released,
as it is while the host waits on a CUDA event) and the fused exchange is a counter per node in shared
memory that every trainer waits on (what the exchange kernel's arrival flags do).  Round time minus the
emulated device time is what the Python control plane costs per round - the quantity that bounds weak
scaling of ``bench.py`` at 8 GPUs and the rounds/s of the 64-node configuration.

    python bench/control_plane_bench.py --nodes 8 --rounds 60 --device-ms 2.6            # bench.py at 8 GPUs
    python bench/control_plane_bench.py --nodes 32 --per-proc 4 --topology line --train-set 8   # config 4, scaled down

Prints one JSON line: ms per round (max over the processes' first nodes), the overhead over the
emulated device time, messages handled per round and node, wait-time percentiles, and the steal time of
the run (a virtual machine that loses its CPUs makes every number here meaningless: keep runs with low steal).
"""

from __future__ import annotations

import argparse
import json
import multiprocessing as mp
import os
import sys
import time
from typing import Any, Dict, List, Optional, Tuple

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _worker(proc: int, n: int, args: Any, counters: Any, gate: Any, out: Any) -> None:
    import threading

    import torch

    torch.set_num_threads(1)
    from p2pfl_b200.communication.sockets import SocketCommunicationProtocol
    from p2pfl_b200.communication.sockets.socket_communication_protocol import SocketTransport
    from p2pfl_b200.learning.learner import NodeLearner
    from p2pfl_b200.management.logger import logger
    from p2pfl_b200.node import Node
    from p2pfl_b200.settings import Settings
    from p2pfl_b200.stages.base_node.vote_train_set_stage import VoteTrainSetStage
    from p2pfl_b200.utils import wait_convergence

    logger.set_level("WARNING")
    Settings.LOG_LEVEL = "WARNING"
    Settings.TRAIN_SET_SIZE = args.train_set or n
    Settings.TTL = max(10, n)
    if args.no_advertise:  # A/B: neighbour lists travel with the periodic beat only (the behaviour before the fix)
        from p2pfl_b200.communication.protocol import BaseProtocol

        BaseProtocol._advertise_direct = lambda self: None  # type: ignore[method-assign]
    if args.no_local_shortcut:  # A/B: every hop over the loopback socket, also between nodes of one process
        Settings.SOCKET_LOCAL_SHORTCUT = False
    if args.flood_status:  # A/B: models_ready / model_initialized flooded network-wide (the reference's way)
        Settings.DIRECT_ROUND_STATUS = False
    Settings.GOSSIP_MODELS_PERIOD = 0.05
    Settings.GOSSIP_PERIOD = 0.0
    Settings.WAIT_HEARTBEATS_CONVERGENCE = 0.2
    Settings.HEARTBEAT_PERIOD = 1.0
    Settings.HEARTBEAT_TIMEOUT = 30.0
    Settings.GRPC_TIMEOUT = 30.0
    Settings.AGGREGATION_TIMEOUT = 60.0
    Settings.VOTE_TIMEOUT = 60.0
    device_s = args.device_ms / 1e3
    port0, k = args.port, args.per_proc
    mine = list(range(proc * k, min(n, proc * k + k)))
    waits: Dict[str, List[float]] = {"vote": [], "finish": [], "handle": [], "deliver_many": []}
    counts: Dict[str, int] = {}

    def index_of(addr: str) -> int:
        return int(addr.rsplit(":", 1)[1]) - port0

    class EmuFused:
        can_prelaunch = True

        def __init__(self, state: Any, train_set: List[str]) -> None:
            self.state, self.train_set = state, train_set
            self.me = index_of(state.addr)
            self.members = [index_of(a) for a in train_set]
            self.seq = 0

        def launch(self, seq: int) -> None:
            self.seq = seq - self.state.seq_base
            counters[self.me] = self.seq  # "my tiles are on their way"

        def finish(self) -> Optional[Tuple[List[str], int]]:
            t = time.perf_counter()
            while min(counters[m] for m in self.members) < self.seq:  # the kernel's wait for its peers' arrival flags
                time.sleep(args.poll_us * 1e-6)
            if self.me == mine[0]:
                waits["finish"].append((time.perf_counter() - t) * 1e3)
            return list(self.train_set), 100 * len(self.train_set)

        def run(self, seq: int) -> Optional[Tuple[List[str], int]]:
            self.launch(seq)
            return self.finish()

    class EmuProtocol(SocketCommunicationProtocol):
        def on_message(self, msg: Any, via: Any = None) -> Any:
            first = self.addr.endswith(f":{port0 + mine[0]}")
            if first:
                key = msg.cmd if via is None else "in:" + msg.cmd
                counts[key] = counts.get(key, 0) + 1
            if via is not None or not first:
                return super().on_message(msg, via)
            t = time.perf_counter()
            try:
                return super().on_message(msg, via)
            finally:
                waits["handle"].append((time.perf_counter() - t) * 1e3)

        def plan_fused_exchange(self, state: Any, learner: Any) -> Any:
            if len(state.train_set) < 2 or args.no_fused:
                return None
            return EmuFused(state, sorted(state.train_set))

    class EmuLearner(NodeLearner):
        arena = object()  # "has a device arena": the stage takes the pipelined branch

        def __init__(self, model: Any = None, data: Any = None, self_addr: str = "", epochs: int = 1, **_: Any) -> None:
            self.flat = torch.zeros(16)
            self.addr = self_addr

        def set_model(self, model: Any) -> None: ...
        def set_data(self, data: Any) -> None: ...
        def set_addr(self, addr: str) -> None:
            self.addr = addr

        def encode_parameters(self, params: Any = None) -> Any:
            import pickle

            return pickle.dumps([self.flat.numpy()])

        def decode_parameters(self, data: Any) -> Any:
            import pickle

            return {"w": torch.from_numpy(pickle.loads(data)[0])}

        def check_parameters(self, params: Any) -> None: ...
        def set_parameters(self, params: Any) -> None: ...
        def get_parameters(self) -> Dict[str, Any]:
            return {"w": self.flat}

        def live_params(self) -> Any:
            return self.flat

        def set_epochs(self, epochs: int) -> None: ...
        def interrupt_fit(self) -> None: ...
        def get_num_samples(self) -> Tuple[int, int]:
            return 100, 10

        def evaluate(self) -> Dict[str, float]:
            return {"test_loss": 1.0, "test_metric": 0.5}

        def evaluate_async(self) -> Any:
            return 1

        def evaluate_finish(self, handle: Any) -> Dict[str, float]:
            return self.evaluate()

        def fit(self, after_launch: Any = None) -> None:
            if after_launch is not None:
                after_launch()
            time.sleep(device_s)  # the host waits for the validation result of the epoch

    orig_many = SocketTransport.deliver_many

    def timed_many(self: Any, targets: Any, env: Any) -> Any:
        t = time.perf_counter()
        try:
            return orig_many(self, targets, env)
        finally:
            waits["deliver_many"].append((time.perf_counter() - t) * 1e3)

    SocketTransport.deliver_many = timed_many  # type: ignore[method-assign]
    orig_agg = VoteTrainSetStage._aggregate_votes

    def timed_agg(ctx: Any) -> Any:
        t = time.perf_counter()
        try:
            return orig_agg(ctx)
        finally:
            if ctx.state.addr.endswith(f":{port0 + mine[0]}"):
                waits["vote"].append((time.perf_counter() - t) * 1e3)

    VoteTrainSetStage._aggregate_votes = staticmethod(timed_agg)  # type: ignore[method-assign]

    timeline: List[Any] = []
    if args.timeline:  # stage entries / exits and model sends of this process's first node, for the last rounds
        from p2pfl_b200.stages.base_node import gossip_model_stage, round_finished_stage, train_stage, wait_agg_models_stage

        for cls in (train_stage.TrainStage, wait_agg_models_stage.WaitAggregatedModelsStage, gossip_model_stage.GossipModelStage,
                    round_finished_stage.RoundFinishedStage, VoteTrainSetStage):  # fmt: skip
            def wrap(cls: Any = cls) -> None:
                inner = cls.execute

                def execute(ctx: Any) -> Any:
                    t = time.time()
                    try:
                        return inner(ctx)
                    finally:
                        if ctx.state.addr.endswith(f":{port0 + mine[0]}"):
                            timeline.append((round(t * 1e3 % 1e6, 1), index_of(ctx.state.addr), ctx.state.round, cls.__name__, round((time.time() - t) * 1e3, 1)))

                cls.execute = staticmethod(execute)

            wrap()
        orig_on_weights = EmuProtocol.on_weights

        def on_weights(self: Any, w: Any) -> Any:
            if self.addr.endswith(f":{port0 + mine[0]}"):
                timeline.append((round(time.time() * 1e3 % 1e6, 1), index_of(self.addr), w.round, f"<- {w.cmd} from {index_of(w.source)}", 0))
            return orig_on_weights(self, w)

        EmuProtocol.on_weights = on_weights  # type: ignore[method-assign]
    nodes = [Node(None, None, address=f"127.0.0.1:{port0 + i}", learner=EmuLearner, protocol=EmuProtocol) for i in mine]
    for nd in nodes:
        nd.start()
    gate.wait()
    for i, nd in zip(mine, nodes):
        peers = list(range(i)) if args.topology == "full" else ([i - 1] if i > 0 else [])
        if args.topology == "ring" and i == n - 1 and n > 2:
            peers.append(0)
        for r in peers:
            nd.connect(f"127.0.0.1:{port0 + r}")
    gate.wait()
    wait_convergence(nodes, n - 1, wait=60, only_direct=args.topology == "full")
    time.sleep(args.settle)
    gate.wait()

    W, K = args.warmup, args.rounds
    marks: Dict[str, Any] = {}
    first = nodes[0]

    def on_round(r: int) -> None:
        if r == W:
            marks["t0"] = time.perf_counter()
            marks["tim0"] = dict(first.live_timings)
            marks["c0"] = dict(counts)
        elif r == W + K:
            marks["t1"] = time.perf_counter()
            marks["tim1"] = dict(first.live_timings)
            marks["c1"] = dict(counts)

    def dump_if_stuck() -> None:
        time.sleep(args.stuck_after)
        if "t1" in marks:
            return
        import traceback

        lines = []
        for nd in nodes:
            st = nd.state
            lines.append(f"[node {index_of(st.addr)}] STUCK: round {st.round} status {st.status} train_set {len(st.train_set)} "
                         f"votes { {r: len(v) for r, v in st.train_set_votes.items()} } counters {list(counters[:n])} "
                         f"neighbors {len(nd._communication_protocol.get_neighbors())}")  # fmt: skip
        for tid, frame in sys._current_frames().items():
            name = next((t.name for t in threading.enumerate() if t.ident == tid), str(tid))
            if "learning" in name.lower():
                lines.append(f"--- {name}\n" + "".join(traceback.format_stack(frame)[-5:]))
        sys.stderr.write("\n".join(lines) + "\n")
        sys.stderr.flush()

    threading.Thread(target=dump_if_stuck, daemon=True).start()
    first.state.round_listeners = [on_round]
    if proc == 0:
        first.set_start_learning(rounds=W + K + 1, epochs=1)
    deadline = time.time() + args.stuck_after + 60
    while first.state.finished.is_set() and time.time() < deadline:
        time.sleep(0.001)
    for nd in nodes:
        nd.wait_learning_finished(max(1.0, deadline - time.time()))
    res: Dict[str, Any] = {"proc": proc}
    if "t1" in marks:
        res["ms_per_round"] = (marks["t1"] - marks["t0"]) * 1e3 / K
        t0, t1 = marks["tim0"], marks["tim1"]
        res["stages"] = {k_: round((v - t0.get(k_, 0.0)) * 1e3 / K, 3) for k_, v in t1.items()}
        res["msgs_per_round"] = {k_: round((v - marks["c0"].get(k_, 0)) / K, 2) for k_, v in marks["c1"].items()}
        for k_, v in waits.items():
            v = sorted(v[-(K if k_ in ("vote", "finish") else 7 * K):])
            if v:
                res[f"wait_{k_}_ms"] = {"p50": round(v[len(v) // 2], 3), "p90": round(v[int(len(v) * 0.9)], 3), "max": round(v[-1], 3)}
    from p2pfl_b200.utils import lockcheck

    if lockcheck.enabled():  # P2PFL_LOCKCHECK=1: cycles in the lock-order graph seen during the run
        res["lock_order_violations"] = [list(v) for v in lockcheck.violations()]
        res["lock_order_edges"] = sum(len(v) for v in lockcheck.edges().values())
    if timeline:
        res["timeline"] = timeline[-60:]
    out.put(res)
    for nd in nodes:
        nd.stop()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=8)
    ap.add_argument("--per-proc", type=int, default=1, help="nodes per process (config 4: 8 per GPU)")
    ap.add_argument("--topology", default="full", choices=["full", "line", "ring"])
    ap.add_argument("--rounds", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--device-ms", type=float, default=2.6)
    ap.add_argument("--port", type=int, default=21200)
    ap.add_argument("--settle", type=float, default=0.0, help="seconds between the last connect and the start of learning")
    ap.add_argument("--poll-us", type=float, default=30.0, help="sleep between two polls of the emulated exchange")
    ap.add_argument("--stuck-after", type=float, default=60.0, help="dump every node's state if the run has not finished by then")
    ap.add_argument("--timeline", action="store_true", help="print the last stage transitions of every process's first node")
    ap.add_argument("--no-fused", action="store_true", help="exchange models through add_model messages (the host path)")
    ap.add_argument("--train-set", type=int, default=0, help="TRAIN_SET_SIZE (default: every node trains)")
    ap.add_argument("--no-advertise", action="store_true", help="A/B: do not advertise direct neighbours on change")
    ap.add_argument("--no-local-shortcut", action="store_true", help="A/B: nodes of one process also talk over sockets")
    ap.add_argument("--flood-status", action="store_true", help="A/B: flood models_ready / model_initialized network-wide")
    args = ap.parse_args()
    ctx = mp.get_context("spawn")
    n = args.nodes
    n_proc = (n + args.per_proc - 1) // args.per_proc

    def steal() -> int:
        try:
            with open("/proc/stat") as fh:
                return int(fh.readline().split()[8])
        except Exception:
            return -1

    steal0 = steal()
    counters = ctx.Array("q", n, lock=False)
    gate = ctx.Barrier(n_proc)
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(p, n, args, counters, gate, out)) for p in range(n_proc)]
    for p in procs:
        p.start()
    res = sorted((out.get(timeout=args.stuck_after + 240) for _ in procs), key=lambda r: r["proc"])
    for p in procs:
        p.join(30)
    ms = max(r.get("ms_per_round", float("nan")) for r in res)
    r0 = res[0]
    line = {
        "bench": "control_plane", "nodes": n, "per_proc": args.per_proc, "topology": args.topology,
        "train_set": args.train_set or n, "fused": not args.no_fused, "rounds": args.rounds, "device_ms_emulated": args.device_ms,
        "ms_per_round": round(ms, 3), "host_overhead_ms": round(ms - args.device_ms, 3),
        "stages_node0": r0.get("stages"), "msgs_per_round_node0": r0.get("msgs_per_round"),
        "node0_ms": {k: v for k, v in r0.items() if k.startswith("wait_")},
        "cpus": os.cpu_count(), "steal_jiffies_whole_run": steal() - steal0,
        "advertise_on_change": not args.no_advertise, "direct_round_status": not args.flood_status, "local_shortcut": not args.no_local_shortcut,
    }  # fmt: skip
    if any("lock_order_violations" in r for r in res):
        line["lock_order_violations"] = [v for r in res for v in r.get("lock_order_violations", [])]
        line["lock_order_edges"] = max(r.get("lock_order_edges", 0) for r in res)
    print(json.dumps(line), flush=True)
    if args.timeline:
        for e in sorted((e for r in res for e in r.get("timeline", [])), key=lambda e: e[0]):
            print(e)


if __name__ == "__main__":
    main()
