#!/usr/bin/env python
"""Headline benchmark: federated rounds of the MNIST-shaped MLP (BASELINE.json config 2).

    python bench.py --gpus N --steps K --warmup W            # this framework
    python bench.py --impl reference --gpus N --steps K ...  # unmodified reference (baseline/_ref)

One federated node per GPU (one process per GPU under torchrun), fully connected,
FedAvg, every node in the train set, 1 local epoch per round over a fixed
per-node shard (weak scaling: 60000/8 MNIST-shaped samples per node whatever N
is).  A "step" is one federated ROUND: evaluate -> local epoch (+ validation)
-> model exchange + FedAvg -> diffusion bookkeeping.  W warm-up rounds, then
exactly K timed rounds bracketed by barrier + cuda synchronize, timed with CUDA
events, max over ranks.  `value` = training samples/s of the whole job.

Prints ONE JSON line (rank 0).  See DESIGN.md "bench contract" for the fields.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from typing import Any, Dict, List, Optional

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SAMPLES_POOL = 60000
TEST_POOL = 10000
PARTS = 8  # shard size is fixed at pool/8 per node (weak scaling)
BATCH = 32
L2_FLUSH_BYTES = 256 << 20


# --------------------------------------------------------------------- helpers
def dist_env() -> Dict[str, int]:
    return {
        "rank": int(os.environ.get("RANK", "0")),
        "local_rank": int(os.environ.get("LOCAL_RANK", "0")),
        "world": int(os.environ.get("WORLD_SIZE", "1")),
    }


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = (
        "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
        "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    )

    def __init__(self, gpu_index: int) -> None:
        self.gpu = gpu_index
        self.proc: Optional[subprocess.Popen] = None
        self.lines: List[str] = []
        self._t: Optional[threading.Thread] = None

    def start(self) -> None:
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.gpu)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True,
            )  # fmt: skip
        except Exception:
            self.proc = None
            return

        def pump() -> None:
            assert self.proc is not None and self.proc.stdout is not None
            for line in self.proc.stdout:
                self.lines.append(line.strip())

        self._t = threading.Thread(target=pump, daemon=True)
        self._t.start()

    def stop(self) -> Dict[str, Any]:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {
            "sm_mhz": statistics.median(sm) if sm else None,
            "sm_max_mhz": max(mx) if mx else None,
            "power_w_max": max(power) if power else None,
            "samples": len(sm),
            "reasons": sorted(reasons),
        }


# ------------------------------------------------------------------ own arm
def run_ours(args: argparse.Namespace) -> Dict[str, Any]:
    import torch
    import torch.distributed as dist

    import __graft_entry__ as entry

    env = dist_env()
    rank, local, world = env["rank"], env["local_rank"], env["world"]
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device")
    torch.cuda.set_device(local)
    if rank == 0:
        entry.build()
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
        dist.barrier()

    from p2pfl_b200 import ops
    from p2pfl_b200.learning.b200_learner import B200Learner
    from p2pfl_b200.learning.data import MnistFederatedDM
    from p2pfl_b200.management.logger import logger
    from p2pfl_b200.models import MLP
    from p2pfl_b200.node import Node
    from p2pfl_b200.settings import Settings
    from p2pfl_b200.utils import wait_convergence

    ops.lib()
    logger.set_level("WARNING")
    Settings.LOG_LEVEL = "WARNING"
    Settings.TRAIN_SET_SIZE = world
    Settings.DEVICE_INBOX_SLOTS = max(world, 2)
    Settings.GOSSIP_MODELS_PERIOD = 0.05
    Settings.GOSSIP_PERIOD = 0.0
    Settings.WAIT_HEARTBEATS_CONVERGENCE = 0.2
    Settings.HEARTBEAT_PERIOD = 1.0
    Settings.HEARTBEAT_TIMEOUT = 30.0
    Settings.GRPC_TIMEOUT = 30.0
    Settings.AGGREGATION_TIMEOUT = 120.0
    Settings.VOTE_TIMEOUT = 120.0

    if world > 1:
        from p2pfl_b200.communication.nvlink import NvlinkCommunicationProtocol as Proto

        port0 = int(os.environ.get("MASTER_PORT", "29500")) + 211
        address = f"127.0.0.1:{port0 + rank}"
    else:
        from p2pfl_b200.communication.memory import InMemoryCommunicationProtocol as Proto

        address = "bench-node-0"

    dm = MnistFederatedDM(rank % PARTS, PARTS, batch_size=BATCH, n_train=SAMPLES_POOL, n_test=TEST_POOL, seed=1234 + rank // PARTS)
    n_train, _ = dm.num_samples()
    node = Node(MLP(seed=rank), dm, address=address, learner=B200Learner, protocol=Proto, device=f"cuda:{local}")
    node.start()
    if world > 1:
        dist.barrier()
        for r in range(rank):
            node.connect(f"127.0.0.1:{port0 + r}")
        dist.barrier()
        wait_convergence([node], world - 1, wait=30, only_direct=True)
        dist.barrier()

    flush = torch.empty(L2_FLUSH_BYTES // 4, dtype=torch.float32, device=f"cuda:{local}")
    W, K = args.warmup, args.steps
    total_rounds = W + K + 1  # one untimed cool-down round so the K-th boundary is an ordinary one

    def measure(data_mode: str) -> Dict[str, Any]:
        node.learner_kwargs["data_mode"] = data_mode
        marks: Dict[str, Any] = {}
        sampler = ClockSampler(local)

        def on_round(r: int) -> None:
            # r = number of rounds completed on this node
            with torch.cuda.stream(node.state.learner.stream):
                flush.zero_()  # L2 flush between rounds (inputs are smaller than the 126 MB L2)
            if r == W:
                torch.cuda.synchronize()
                if world > 1:
                    dist.barrier()
                torch.cuda.synchronize()
                if rank == 0:
                    sampler.start()
                marks["launch0"] = ops.launch_count()
                lr = node.state.learner
                marks["h2d0"], marks["d2h0"] = lr.h2d_bytes, lr.d2h_bytes
                marks["tim0"] = dict(node.live_timings)
                marks["ht0"] = dict(getattr(node.state.learner, "host_seconds", {}))
                marks["t0"] = time.perf_counter()
                marks["e0"] = torch.cuda.Event(enable_timing=True)
                marks["e0"].record(torch.cuda.current_stream())
            elif r == W + K:
                lr = node.state.learner
                lr.stream.synchronize()
                marks["e1"] = torch.cuda.Event(enable_timing=True)
                marks["e1"].record(torch.cuda.current_stream())
                torch.cuda.synchronize()
                marks["t1"] = time.perf_counter()
                marks["tim1"] = dict(node.live_timings)
                marks["ht1"] = dict(getattr(lr, "host_seconds", {}))
                marks["launch1"] = ops.launch_count()
                marks["h2d1"], marks["d2h1"] = lr.h2d_bytes, lr.d2h_bytes
                if world > 1:
                    dist.barrier()
                if rank == 0:
                    marks["clocks"] = sampler.stop()

        node.state.round_listeners = [on_round]
        if world > 1:
            dist.barrier()
        if rank == 0:
            node.set_start_learning(rounds=total_rounds, epochs=1)
        deadline = time.time() + args.timeout
        while node.state.finished.is_set() and time.time() < deadline:  # wait for the experiment to begin here
            time.sleep(0.001)
        if not node.wait_learning_finished(max(1.0, deadline - time.time())):
            raise SystemExit(f"rank {rank}: experiment did not finish in {args.timeout}s")
        if "e1" not in marks:
            raise SystemExit(f"rank {rank}: timed region never closed (rounds seen: {sorted(marks)})")
        ms = marks["e0"].elapsed_time(marks["e1"])
        t = torch.tensor([ms, (marks["t1"] - marks["t0"]) * 1e3], dtype=torch.float64, device=f"cuda:{local}")
        cnt = torch.tensor(
            [marks["launch1"] - marks["launch0"], marks["h2d1"] - marks["h2d0"], marks["d2h1"] - marks["d2h0"]],
            dtype=torch.float64, device=f"cuda:{local}",
        )  # fmt: skip
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        trace_dir = os.environ.get("P2PFL_BENCH_TRACE")
        if trace_dir:
            # per-rank host timeline of the timed region (ms per round), for scaling analysis
            os.makedirs(trace_dir, exist_ok=True)
            per = {k: round((v - marks.get("tim0", {}).get(k, 0.0)) * 1e3 / K, 3) for k, v in marks.get("tim1", {}).items()}
            per.update({k: round((v - marks.get("ht0", {}).get(k, 0.0)) * 1e3 / K, 3) for k, v in marks.get("ht1", {}).items()})
            per["device_ms_per_round"] = round(ms / K, 3)
            with open(os.path.join(trace_dir, f"stage_{data_mode}_rank{rank}.json"), "w") as fh:
                json.dump(per, fh)
        return {
            "ms": float(t[0]), "wall_ms": float(t[1]), "launches": int(cnt[0]),
            "h2d": float(cnt[1]), "d2h": float(cnt[2]), "clocks": marks.get("clocks"),
            "timings": {k: v - marks.get("tim0", {}).get(k, 0.0) for k, v in marks.get("tim1", {}).items()},
        }  # fmt: skip

    dev = measure("resident")
    e2e = measure("stream")
    node.stop()

    samples = world * n_train * K
    value = samples / (dev["ms"] / 1e3)
    e2e_value = samples / (e2e["ms"] / 1e3)
    out = {
        "metric": "fl_train_samples_per_sec",
        "value": value,
        "unit": "samples/s",
        "n_gpus": world,
        "steps": K,
        "warmup": W,
        "ms_per_step": dev["ms"] / K,
        "rounds_per_sec": K / (dev["ms"] / 1e3),
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": None,
        "dtype": "bf16",
        "data": "synthetic",
        "impl": "p2pfl_b200",
        "config": {
            "model": "MNIST-shaped MLP 784-256-128-10 (235146 params)",
            "baseline_config": "MNIST-shaped MLP, N nodes (one per B200), fully connected topology, FedAvg",
            "nodes": world,
            "topology": "full",
            "aggregator": "FedAvg, all nodes in the train set",
            "epochs_per_round": 1,
            "train_samples_per_node": n_train,
            "global_batch": BATCH * world,
            "batch_per_node": BATCH,
            "seq_len": None,
            "parallelism": f"fl{world} (one federated node per GPU)",
            "step_definition": "one federated round: evaluate + 1 local epoch (+validation) + model exchange + FedAvg",
            "l2": f"flushed between rounds ({L2_FLUSH_BYTES >> 20} MiB memset)",
            "control_plane": "in-memory (1 node)" if world == 1 else "gRPC loopback + NVLink device data plane",
        },
        "clocks": dev["clocks"],
        "e2e": {
            "value": e2e_value,
            "unit": "samples/s",
            "ms_per_step": e2e["ms"] / K,
            "h2d_bytes_per_step": e2e["h2d"] / K / world,
            "d2h_bytes_per_step": e2e["d2h"] / K / world,
            "clocks": e2e["clocks"],
            "note": "per round: train/val/test splits copied from pinned host memory, loss+metric read back",
        },
        "gpu_launches": dev["launches"],
        "wall_ms_per_step": dev["wall_ms"] / K,
        "stage_ms_per_step_rank0": {k: round(v * 1e3 / K, 3) for k, v in dev["timings"].items()},
    }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out if rank == 0 else {}


# ------------------------------------------------------------- reference arm
def run_reference(args: argparse.Namespace) -> Dict[str, Any]:
    env = dist_env()
    rank, local, world = env["rank"], env["local_rank"], env["world"]
    ref_root = os.path.join(ROOT, "baseline", "_ref")
    shim = os.path.join(ROOT, "baseline", "shim")
    if not os.path.isdir(os.path.join(ref_root, "p2pfl")):
        return {"impl": "reference", "unavailable": "baseline/_ref is missing (run baseline/install_reference.sh)"} if rank == 0 else {}
    import torch
    import torch.distributed as dist

    if world > 1:
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
        dist.barrier()
    out: Dict[str, Any] = {}
    if rank == 0:
        # The reference's own multi-node mode is one process with N Node objects
        # (SURVEY §0.4); its multi-process gRPC mode cannot train (Q8: a remote
        # `metrics` message raises KeyError and evicts the sender).  Ranks > 0 idle.
        sys.path[:0] = [shim, ref_root]
        from baseline.reference_runner import run_reference_experiment

        out = run_reference_experiment(world, args.steps, args.warmup, args.ref_fast, args.timeout)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ref-fast", action="store_true", help="zero the reference's pacing sleeps (not the stock path)")
    ap.add_argument("--timeout", type=float, default=1500.0)
    args = ap.parse_args()
    env = dist_env()
    if args.gpus != env["world"] and env["world"] == 1 and args.gpus > 1:
        # convenience: re-exec under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29512", os.path.abspath(__file__)] + sys.argv[1:]  # fmt: skip
        raise SystemExit(subprocess.call(cmd))
    if args.warmup < 3:
        args.warmup = 3
    res = run_reference(args) if args.impl == "reference" else run_ours(args)
    if env["rank"] == 0:
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
