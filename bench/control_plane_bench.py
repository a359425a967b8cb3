#!/usr/bin/env python
"""Host-side cost of a federated round, measured WITHOUT a GPU.

One process per node (as on the 8-GPU box), the real node / stage machine / socket control plane, and
an emulated device: ``fit`` sleeps for the device time of the round (the GIL is released, as it is
while the host waits on a CUDA event) and the fused exchange is a counter per rank in shared memory
that every rank waits on (what the exchange kernel's arrival flags do).  Round time minus the emulated
device time is what the Python control plane costs per round - the quantity that bounds weak scaling
of ``bench.py`` at 8 GPUs.

    python bench/control_plane_bench.py --nodes 8 --rounds 60 --device-ms 2.6 [--profile]

Prints one JSON line: ms per round (max over ranks), the overhead over the emulated device time and,
with ``--profile``, rank 0's hottest functions (all threads).
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


def _worker(rank: int, n: int, args: Any, port0: int, counters: Any, gate: Any, out: Any) -> None:
    import torch

    torch.set_num_threads(1)
    from p2pfl_b200.communication.sockets import SocketCommunicationProtocol
    from p2pfl_b200.learning.learner import NodeLearner
    from p2pfl_b200.management.logger import logger
    from p2pfl_b200.node import Node
    from p2pfl_b200.settings import Settings
    from p2pfl_b200.utils import wait_convergence

    logger.set_level("WARNING")
    Settings.LOG_LEVEL = "WARNING"
    Settings.TRAIN_SET_SIZE = args.train_set or n
    if args.no_advertise:  # A/B: neighbour lists travel with the periodic beat only (the behaviour before the fix)
        from p2pfl_b200.communication.protocol import BaseProtocol

        BaseProtocol._advertise_direct = lambda self: None  # type: ignore[method-assign]
    Settings.GOSSIP_MODELS_PERIOD = 0.05
    Settings.GOSSIP_PERIOD = 0.0
    Settings.WAIT_HEARTBEATS_CONVERGENCE = 0.2
    Settings.HEARTBEAT_PERIOD = 1.0
    Settings.HEARTBEAT_TIMEOUT = 30.0
    Settings.GRPC_TIMEOUT = 30.0
    Settings.AGGREGATION_TIMEOUT = 60.0
    Settings.VOTE_TIMEOUT = 60.0
    device_s = args.device_ms / 1e3

    class EmuFused:
        can_prelaunch = True

        def __init__(self, state: Any, train_set: List[str]) -> None:
            self.state, self.train_set = state, train_set
            self.members = [int(a.rsplit(":", 1)[1]) - port0 for a in train_set]
            self.seq = 0

        def launch(self, seq: int) -> None:
            self.seq = seq - state_seq_base[0]
            counters[rank] = self.seq  # "my tiles are on their way"

        def finish(self) -> Optional[Tuple[List[str], int]]:
            while min(counters[m] for m in self.members) < self.seq:  # the kernel's wait for every peer's arrival flags
                time.sleep(args.poll_us * 1e-6)
            return list(self.train_set), 100 * len(self.train_set)

        def run(self, seq: int) -> Optional[Tuple[List[str], int]]:
            self.launch(seq)
            return self.finish()

    state_seq_base = [0]
    waits: Dict[str, List[float]] = {"vote": [], "finish": [], "handle": [], "deliver_many": []}

    counts: Dict[str, int] = {}

    class EmuProtocol(SocketCommunicationProtocol):
        def on_message(self, msg: Any, via: Any = None) -> Any:
            k = msg.cmd if via is None else "in:" + msg.cmd
            counts[k] = counts.get(k, 0) + 1
            if via is not None:
                return super().on_message(msg, via)
            t = time.perf_counter()
            try:
                return super().on_message(msg, via)
            finally:
                waits["handle"].append((time.perf_counter() - t) * 1e3)

        def plan_fused_exchange(self, state: Any, learner: Any) -> Any:
            if len(state.train_set) < 2 or args.no_fused:
                return None
            state_seq_base[0] = state.seq_base
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

    from p2pfl_b200.stages.base_node.vote_train_set_stage import VoteTrainSetStage

    from p2pfl_b200.communication.sockets.socket_communication_protocol import SocketTransport

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
            waits["vote"].append((time.perf_counter() - t) * 1e3)

    VoteTrainSetStage._aggregate_votes = staticmethod(timed_agg)  # type: ignore[method-assign]
    orig_finish = EmuFused.finish

    def timed_finish(self: Any) -> Any:
        t = time.perf_counter()
        try:
            return orig_finish(self)
        finally:
            waits["finish"].append((time.perf_counter() - t) * 1e3)

    EmuFused.finish = timed_finish  # type: ignore[method-assign]

    node = Node(None, None, address=f"127.0.0.1:{port0 + rank}", learner=EmuLearner, protocol=EmuProtocol)
    node.start()
    gate.wait()
    for r in range(rank):
        node.connect(f"127.0.0.1:{port0 + r}")
    gate.wait()
    wait_convergence([node], n - 1, wait=30, only_direct=True)
    time.sleep(args.settle)
    gate.wait()

    W, K = args.warmup, args.rounds
    marks: Dict[str, float] = {}
    prof = None

    def on_round(r: int) -> None:
        nonlocal prof
        if r == W:
            marks["t0"] = time.perf_counter()
            marks["tim0"] = dict(node.live_timings)  # type: ignore[assignment]
            marks["c0"] = dict(counts)  # type: ignore[assignment]
            if args.profile and rank == 0:
                import cProfile
                import threading

                prof = cProfile.Profile()
                prof.enable()
                threading.setprofile(lambda *a: None)  # (other threads: see --profile-threads)
        elif r == W + K:
            if prof is not None:
                prof.disable()
            marks["t1"] = time.perf_counter()
            marks["tim1"] = dict(node.live_timings)  # type: ignore[assignment]
            marks["c1"] = dict(counts)  # type: ignore[assignment]

    def dump_if_stuck() -> None:
        time.sleep(args.stuck_after)
        if "t1" in marks:
            return
        import traceback

        st = node.state
        lines = [f"[rank {rank}] STUCK: round {st.round} status {st.status} train_set {len(st.train_set)} "
                 f"votes {{r: sorted(v) for r, v in st.train_set_votes.items()}} = { {r: len(v) for r, v in st.train_set_votes.items()} } "
                 f"counters {list(counters[:n])} neighbors {len(node._communication_protocol.get_neighbors())}"]  # fmt: skip
        for tid, frame in sys._current_frames().items():
            name = next((t.name for t in __import__("threading").enumerate() if t.ident == tid), str(tid))
            if "learning" in name.lower() or "gossiper" in name.lower():
                lines.append(f"--- {name}\n" + "".join(traceback.format_stack(frame)[-6:]))
        sys.stderr.write("\n".join(lines) + "\n")
        sys.stderr.flush()

    import threading as _th

    _th.Thread(target=dump_if_stuck, daemon=True).start()
    node.state.round_listeners = [on_round]
    if rank == 0:
        node.set_start_learning(rounds=W + K + 1, epochs=1)
    deadline = time.time() + 120
    while node.state.finished.is_set() and time.time() < deadline:
        time.sleep(0.001)
    node.wait_learning_finished(120)
    res: Dict[str, Any] = {"rank": rank}
    if "t1" in marks:
        res["ms_per_round"] = (marks["t1"] - marks["t0"]) * 1e3 / K
        t0, t1 = marks["tim0"], marks["tim1"]
        res["stages"] = {k: round((v - t0.get(k, 0.0)) * 1e3 / K, 3) for k, v in t1.items()}  # type: ignore[union-attr]
        res["msgs_per_round"] = {k: round((v - marks["c0"].get(k, 0)) / K, 2) for k, v in marks["c1"].items()}  # type: ignore[union-attr]
        for k, v in waits.items():
            v = sorted(v[-(K if k in ("vote", "finish") else 7 * K):])
            if v:
                res[f"wait_{k}_ms"] = {"p50": round(v[len(v) // 2], 3), "p90": round(v[int(len(v) * 0.9)], 3), "max": round(v[-1], 3)}
    if prof is not None:
        import io
        import pstats

        s = io.StringIO()
        pstats.Stats(prof, stream=s).sort_stats("tottime").print_stats(25)
        res["profile"] = s.getvalue()
    out.put(res)
    node.stop()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=8)
    ap.add_argument("--rounds", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--device-ms", type=float, default=2.6)
    ap.add_argument("--port", type=int, default=41200)
    ap.add_argument("--settle", type=float, default=0.0, help="seconds between the last connect and the start of learning")
    ap.add_argument("--poll-us", type=float, default=30.0, help="sleep between two polls of the emulated exchange")
    ap.add_argument("--stuck-after", type=float, default=40.0, help="dump every rank's state if the run has not finished by then")
    ap.add_argument("--no-fused", action="store_true", help="exchange models through add_model messages (the host path)")
    ap.add_argument("--train-set", type=int, default=0, help="TRAIN_SET_SIZE (default: every node trains)")
    ap.add_argument("--no-advertise", action="store_true", help="A/B: do not advertise direct neighbours on change")
    ap.add_argument("--profile", action="store_true")
    args = ap.parse_args()
    ctx = mp.get_context("spawn")
    n = args.nodes

    def steal() -> int:
        try:
            with open("/proc/stat") as fh:
                return int(fh.readline().split()[8])
        except Exception:
            return -1

    steal0 = steal()
    counters = ctx.Array("q", n, lock=False)
    gate = ctx.Barrier(n)
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, n, args, args.port, counters, gate, out)) for r in range(n)]
    for p in procs:
        p.start()
    res = [out.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(30)
    ms = max(r.get("ms_per_round", float("nan")) for r in res)
    line = {
        "bench": "control_plane", "nodes": n, "rounds": args.rounds, "device_ms_emulated": args.device_ms,
        "ms_per_round": round(ms, 3), "host_overhead_ms": round(ms - args.device_ms, 3),
        "stages_rank0": next(r.get("stages") for r in res if r["rank"] == 0),
        "msgs_per_round_rank0": next(r.get("msgs_per_round") for r in res if r["rank"] == 0),
        "waits": {r["rank"]: [r.get("wait_vote_ms", {}).get("p50"), r.get("wait_vote_ms", {}).get("p90"),
                              r.get("wait_finish_ms", {}).get("p50"), r.get("wait_finish_ms", {}).get("p90")] for r in sorted(res, key=lambda r: r["rank"])},
        "bundles_in_per_rank": [r.get("msgs_per_round", {}).get("bundle") for r in sorted(res, key=lambda r: r["rank"])],
        "rank0_ms": {k: v for r in res if r["rank"] == 0 for k, v in r.items() if k.startswith("wait_")},
        "cpus": os.cpu_count(), "steal_jiffies_whole_run": steal() - steal0, "advertise_on_change": not args.no_advertise,
    }  # fmt: skip
    print(json.dumps(line), flush=True)
    for r in res:
        if "profile" in r:
            print(r["profile"])


if __name__ == "__main__":
    main()
