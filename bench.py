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
import datetime
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
    """SM clock, power and throttle reasons DURING the timed region (B200_PROFILING.md recipe).

    In-process NVML thread, 100 ms period (one per rank: every rank carries the same, negligible
    load - a ``nvidia-smi -lms`` child on rank 0 alone made rank 0 the straggler of every
    exchange).  Falls back to the ``nvidia-smi`` query loop when the NVML binding is missing."""

    Q = (
        "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
        "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    )
    PERIOD = 0.1

    def __init__(self, gpu_index: int) -> None:
        self.gpu = gpu_index
        self.proc: Optional[subprocess.Popen] = None
        self.lines: List[str] = []
        self._t: Optional[threading.Thread] = None
        self._stop = threading.Event()
        self._nvml: Any = None
        self._handle: Any = None
        self._rows: List[Any] = []  # (sm MHz, max MHz, watts, reason bits)

    def _nvml_open(self) -> bool:
        try:
            import pynvml

            pynvml.nvmlInit()
            handle = None
            try:
                import torch

                uuid = str(torch.cuda.get_device_properties(self.gpu).uuid)
                handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
            except Exception:
                handle = pynvml.nvmlDeviceGetHandleByIndex(self.gpu)
            pynvml.nvmlDeviceGetClockInfo(handle, pynvml.NVML_CLOCK_SM)
            self._nvml, self._handle = pynvml, handle
            return True
        except Exception:
            self._nvml = None
            return False

    def start(self) -> None:
        self._stop.clear()
        if self._nvml is not None or self._nvml_open():
            nv, h = self._nvml, self._handle
            mx = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))

            def poll() -> None:
                while True:
                    try:
                        try:
                            bits = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                        except Exception:
                            bits = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                        self._rows.append((float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)), mx,
                                           nv.nvmlDeviceGetPowerUsage(h) / 1000.0, bits))  # fmt: skip
                    except Exception:
                        pass
                    if self._stop.wait(self.PERIOD):
                        return

            self._t = threading.Thread(target=poll, daemon=True)
            self._t.start()
            return
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
        if self._nvml is not None:
            self._stop.set()
            if self._t is not None:
                self._t.join(timeout=1.0)
            nv = self._nvml
            rows, self._rows = self._rows, []
            names = {
                "hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap,
            }
            reasons = sorted(k for k, bit in names.items() if any(r[3] & bit for r in rows))
            return {
                "sm_mhz": statistics.median(r[0] for r in rows) if rows else None,
                "sm_max_mhz": max(r[1] for r in rows) if rows else None,
                "power_w_max": max(r[2] for r in rows) if rows else None,
                "samples": len(rows),
                "reasons": reasons,
                "how": "in-process NVML thread, 100 ms",
            }
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
            "how": "nvidia-smi -lms 100",
        }


def merge_clocks(per_rank: List[Optional[Dict[str, Any]]]) -> Optional[Dict[str, Any]]:
    """One clocks record for the job: lowest median SM clock over the ranks, union of the reasons."""
    rows = [c for c in per_rank if c and c.get("sm_mhz") is not None]
    if not rows:
        return per_rank[0] if per_rank else None
    return {
        "sm_mhz": min(c["sm_mhz"] for c in rows),
        "sm_max_mhz": max(c["sm_max_mhz"] for c in rows),
        "power_w_max": max(c.get("power_w_max") or 0.0 for c in rows),
        "samples": sum(c.get("samples", 0) for c in rows),
        "reasons": sorted({r for c in rows for r in c.get("reasons", [])}),
        "ranks": len(rows),
        "how": rows[0].get("how"),
    }


def bench_config(world: int, n_train: int) -> Dict[str, Any]:
    """The config block BOTH arms print: what is being measured, not how an arm does it."""
    return {
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
    }


def final_metrics(global_logs: Dict[str, Any]) -> Dict[str, Any]:
    """Last ``test_loss`` / ``test_metric`` each node logged, averaged over the nodes of this process."""
    out: Dict[str, Any] = {}
    for metric in ("test_loss", "test_metric"):
        last = []
        for nodes in global_logs.values():
            for series in nodes.values():
                vals = series.get(metric)
                if vals:
                    last.append(max(vals, key=lambda rv: rv[0]))
        if last:
            out[metric] = sum(v for _, v in last) / len(last)
            out["round"] = max(r for r, _ in last)
            out["nodes_reporting"] = len(last)
    return out


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
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"), timeout=datetime.timedelta(hours=2))
        dist.barrier()

    from p2pfl_b200 import ops
    from p2pfl_b200.learning.b200_learner import B200Learner
    from p2pfl_b200.learning.data import MnistFederatedDM
    from p2pfl_b200 import ops as _ops
    from p2pfl_b200.management.logger import logger

    _ops.set_home_device(local)  # one process per GPU: every thread of this process works on it
    _ops.thread_uses_cuda()
    from p2pfl_b200.models import MLP
    from p2pfl_b200.node import Node
    from p2pfl_b200.settings import Settings
    from p2pfl_b200.utils import wait_convergence

    ops.lib()
    fp32 = args.dtype == "fp32"
    if fp32:
        # precision-parity arm: the eager PyTorch learner of this framework, strict fp32 like the
        # reference (its Lightning loop, TF32 off) - same control plane, same round, library GEMMs
        if args.eager:
            from p2pfl_b200.learning.torch_learner import TorchLearner as Learner
        else:  # same operators and precision, one CUDA graph replay per step instead of ~30 launches
            from p2pfl_b200.learning.graphed_learner import GraphedTorchLearner as Learner

        torch.backends.cuda.matmul.allow_tf32 = False
        torch.backends.cudnn.allow_tf32 = False
    else:
        Learner = B200Learner
    level = os.environ.get("P2PFL_BENCH_LOG", "WARNING")  # DEBUG / INFO while chasing a problem
    logger.set_level(level)
    Settings.LOG_LEVEL = level
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
        if args.control == "grpc":
            from p2pfl_b200.communication.nvlink import NvlinkCommunicationProtocol as Proto
        else:
            from p2pfl_b200.communication.nvlink import NvlinkSocketProtocol as Proto

        if args.exchange == "nccl":
            # baseline arm: same node / stages / learner / control plane, the round's exchange
            # done by NCCL all-gather + average instead of the fused push+FedAvg kernel
            from p2pfl_b200.communication.nvlink.collective_exchange import with_collective_exchange

            Proto = with_collective_exchange(Proto)
        port0 = int(os.environ.get("MASTER_PORT", "29500")) + 211
        address = f"127.0.0.1:{port0 + rank}"
    else:
        from p2pfl_b200.communication.memory import InMemoryCommunicationProtocol as Proto

        address = "bench-node-0"

    dm = MnistFederatedDM(rank % PARTS, PARTS, batch_size=BATCH, n_train=SAMPLES_POOL, n_test=TEST_POOL, seed=1234 + rank // PARTS)
    n_train, _ = dm.num_samples()
    node = Node(MLP(seed=rank), dm, address=address, learner=Learner, protocol=Proto, device=f"cuda:{local}")
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

    def stream_of(lr: Any) -> Any:
        # the fused learner computes on a stream of its own, the eager one on its thread's current stream
        return getattr(lr, "stream", None) or torch.cuda.current_stream()

    def measure(data_mode: str) -> Dict[str, Any]:
        node.learner_kwargs["data_mode"] = data_mode  # (ignored by learners without the parameter)
        marks: Dict[str, Any] = {}
        sampler = ClockSampler(local)

        def on_round(r: int) -> None:
            # r = number of rounds completed on this node
            with torch.cuda.stream(stream_of(node.state.learner)):
                flush.zero_()  # L2 flush between rounds (inputs are smaller than the 126 MB L2)
            if r == W:
                torch.cuda.synchronize()
                # before the barrier: starting the sampler (NVML initialisation, a thread) takes 5-15 ms and
                # differs per rank; started after it, the fastest rank's timed region began with a wait for
                # the slowest rank's sampler (rank intervals of 2.98-3.81 ms per round in one 8-GPU run)
                sampler.start()
                if world > 1:
                    dist.barrier()
                torch.cuda.synchronize()
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
                stream_of(lr).synchronize()
                marks["e1"] = torch.cuda.Event(enable_timing=True)
                marks["e1"].record(torch.cuda.current_stream())
                torch.cuda.synchronize()
                marks["t1"] = time.perf_counter()
                marks["tim1"] = dict(node.live_timings)
                marks["ht1"] = dict(getattr(lr, "host_seconds", {}))
                marks["launch1"] = ops.launch_count()
                marks["h2d1"], marks["d2h1"] = lr.h2d_bytes, lr.d2h_bytes
                marks["clocks"] = sampler.stop()
                if world > 1:
                    dist.barrier()

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
        clocks = [marks.get("clocks")]
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
            clocks = [None] * world
            dist.all_gather_object(clocks, marks.get("clocks"))
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
            "h2d": float(cnt[1]), "d2h": float(cnt[2]), "clocks": merge_clocks(clocks),
            "timings": {k: v - marks.get("tim0", {}).get(k, 0.0) for k, v in marks.get("tim1", {}).items()},
        }  # fmt: skip

    dev = measure("resident")
    # the eager learner has one data path: every batch comes from host memory inside the timed region
    e2e = dev if (fp32 and args.eager) else measure("stream")

    # Correctness of what was just timed: after the last round every node holds the same aggregate
    # (all nodes are trainers; the fused exchange sums in train-set order => identical bits), so a
    # checksum of the live parameters must agree across ranks; and the model must have learned.
    lr = node.state.learner
    with torch.cuda.stream(stream_of(lr)):
        flat = lr.flat.double()
        chk = torch.stack([flat.sum(), flat.abs().sum(), (flat * flat).sum()])
    stream_of(lr).synchronize()
    lo, hi = chk.clone(), chk.clone()
    if world > 1:
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    bit_identical = bool(torch.equal(lo, hi))
    ranks_equal = bool(((hi - lo).abs() <= 1e-6 * hi.abs().clamp_min(1e-30)).all()) and bool(torch.isfinite(chk).all())
    fm = final_metrics(logger.get_global_logs())
    fm_t = torch.tensor([fm.get("test_loss", float("nan")), fm.get("test_metric", float("nan"))], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(fm_t, op=dist.ReduceOp.SUM)
        fm_t /= world
    stats = dict(getattr(node._communication_protocol, "stats", {}))
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
        "dtype": args.dtype,
        "data": "synthetic",
        "impl": "p2pfl_b200",
        "config": bench_config(world, n_train),
        "arm": {
            "control_plane": "in-memory (1 node)" if world == 1 else f"{args.control} loopback + NVLink device data plane",
            "timing": "CUDA events on the learner's stream, max over ranks",
            "exchange": "NCCL all-gather + average (library baseline)" if (world > 1 and args.exchange == "nccl")
                        else "fused push+FedAvg kernel over NVLink peer memory" if world > 1 else "none (1 node)",
            "learner": ("TorchLearner: eager PyTorch" if args.eager else "GraphedTorchLearner: PyTorch operators, one CUDA-graph replay per step")
                       + ", strict fp32 (TF32 off), library GEMMs - the precision-parity arm; gpu_launches counts this library's kernels only" if fp32
                       else "B200Learner: whole-epoch tcgen05 cluster kernel, bf16 operands, fp32 accumulate / master weights / Adam",
        },
        "final": {"test_loss": float(fm_t[0]), "test_accuracy": float(fm_t[1]), "round": fm.get("round"),
                  "param_checksum": [float(v) for v in chk], "ranks_equal": ranks_equal, "bit_identical": bit_identical},
        "data_plane": {k: (round(v, 3) if isinstance(v, float) else v) for k, v in stats.items()},
        "clocks": dev["clocks"],
        "e2e": {
            "value": e2e_value,
            "unit": "samples/s",
            "ms_per_step": e2e["ms"] / K,
            "h2d_bytes_per_step": e2e["h2d"] / K / world,
            "d2h_bytes_per_step": e2e["d2h"] / K / world,
            "clocks": e2e["clocks"],
            "note": "every batch copied from host memory, loss+metric read back" if (fp32 and args.eager)
                    else "per round: train/val/test splits copied from pinned host memory, loss+metric read back",
        },
        "gpu_launches": dev["launches"],
        "wall_ms_per_step": dev["wall_ms"] / K,
        "stage_ms_per_step_rank0": {k: round(v * 1e3 / K, 3) for k, v in dev["timings"].items()},
    }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if not ranks_equal:
        sys.stderr.write(f"rank {rank}: parameters differ across ranks after the last round: min {lo.tolist()} max {hi.tolist()}\n")
        raise SystemExit(3)
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

    host_group = None
    if world > 1:
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"), timeout=datetime.timedelta(hours=2))
        dist.barrier()
        torch.cuda.synchronize()
        # Ranks > 0 idle while rank 0 runs the reference (below).  They must wait on the HOST: an NCCL
        # barrier parks a spinning kernel on every other GPU, which (a) time-slices against the
        # reference's own nodes on those GPUs and (b) trips the NCCL watchdog after 10 minutes.
        host_group = dist.new_group(backend="gloo", timeout=datetime.timedelta(hours=2))
    out: Dict[str, Any] = {}
    if rank == 0:
        # The reference's own multi-node mode is one process with N Node objects
        # (SURVEY §0.4); its multi-process gRPC mode cannot train (Q8: a remote
        # `metrics` message raises KeyError and evicts the sender).  Ranks > 0 idle.
        sys.path[:0] = [shim, ref_root]
        from baseline.reference_runner import run_reference_experiment

        try:
            out = run_reference_experiment(world, args.steps, args.warmup, args.ref_fast, args.timeout, args.ref_budget)
        except Exception as e:  # the line below must still be printed, and the other ranks released
            out = {"impl": "reference", "unavailable": f"reference run failed: {type(e).__name__}: {e}"[:300]}
    if world > 1:
        dist.barrier(group=host_group)
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
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"],
                    help="own arm: bf16 = the fused learner (product path); fp32 = the eager fp32 learner (precision parity with the reference)")
    ap.add_argument("--control", default="socket", choices=["socket", "grpc"],
                    help="control-plane transport between the ranks (the data plane is NVLink either way)")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="own arm, N>1: fused = the product's exchange kernel; nccl = all-gather + average baseline")
    ap.add_argument("--eager", action="store_true", help="with --dtype fp32: the plain eager learner instead of the graphed one")
    ap.add_argument("--ref-budget", type=float, default=600.0,
                    help="reference arm: seconds after which the timed region is closed at the last finished round")
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
