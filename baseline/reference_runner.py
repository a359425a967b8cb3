"""Runs the UNMODIFIED reference (baseline/_ref, Angel3245/p2pfl v0.3.0) on the
benchmark config through its own public API:

    Node(MLP(), datamodule, protocol=InMemoryCommunicationProtocol)
    node.start(); node.connect(...); node.set_start_learning(rounds, epochs)

Only two things are not the reference's own: the Lightning/torchmetrics shim
(baseline/shim, because neither package exists offline) and the synthetic
MNIST-shaped datamodule (the reference's downloads MNIST).  Round boundaries
are observed from outside by polling ``node.state.round``; nothing inside the
reference is patched.  ``Settings.TRAIN_SET_SIZE`` is raised to the node count
(the same override this framework's arm uses) so that every node trains.
"""

from __future__ import annotations

import threading
import time
from typing import Any, Dict, List

import torch


def run_reference_experiment(n_nodes: int, steps: int, warmup: int, fast: bool, timeout: float,
                             budget: float = 600.0) -> Dict[str, Any]:  # fmt: skip
    """``budget``: seconds of run time after which the timed region is closed at the last round
    boundary seen (the line then says ``"truncated": true`` and reports the rounds it did time), so a
    slow stock run still yields a number inside the driver's per-run limit."""
    from p2pfl.communication.memory.memory_communication_protocol import InMemoryCommunicationProtocol
    from p2pfl.learning.pytorch.mnist_examples.models.mlp import MLP
    from p2pfl.management.logger import logger as ref_logger
    from p2pfl.node import Node
    from p2pfl.settings import Settings as RefSettings

    from p2pfl_b200.learning.data import MnistFederatedDM

    import bench

    try:
        ref_logger.set_level("WARNING")
    except Exception:
        pass
    RefSettings.TRAIN_SET_SIZE = max(n_nodes, 1)
    if fast:
        RefSettings.GOSSIP_MODELS_PERIOD = 0
        RefSettings.GOSSIP_PERIOD = 0
        RefSettings.WAIT_HEARTBEATS_CONVERGENCE = 0

    n_gpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    nodes: List[Any] = []
    n_train = 0
    for i in range(n_nodes):
        dm = MnistFederatedDM(i % bench.PARTS, bench.PARTS, batch_size=bench.BATCH, n_train=bench.SAMPLES_POOL,
                              n_test=bench.TEST_POOL, seed=1234 + i // bench.PARTS)  # fmt: skip
        if n_gpu:
            dm.device_hint = f"cuda:{i % n_gpu}"  # read by the shim Trainer: one GPU per node
        n_train = dm.num_samples()[0]
        node = Node(MLP(), dm, protocol=InMemoryCommunicationProtocol, address=f"ref-node-{i}")
        node.start()
        nodes.append(node)
    for i in range(1, n_nodes):
        for j in range(i):
            nodes[i].connect(nodes[j].addr)
    deadline = time.time() + 30
    while any(len(n.get_neighbors(only_direct=True)) != n_nodes - 1 for n in nodes) and time.time() < deadline:
        time.sleep(0.05)

    W, K = warmup, steps
    total = W + K + 1
    marks: Dict[str, Any] = {}
    stop = threading.Event()
    import pytorch_lightning as shim_pl  # the shim counts the bytes it copies (see its COUNTERS)

    used = sorted({i % n_gpu for i in range(n_nodes)}) if n_gpu else []
    flush = {d: torch.empty(bench.L2_FLUSH_BYTES // 4, dtype=torch.float32, device=f"cuda:{d}") for d in used}

    def sync_all() -> None:
        for d in used:
            torch.cuda.synchronize(d)

    def watch() -> None:
        # all nodes finished >= r rounds  <=>  min over nodes of state.round >= r.  A reference node
        # can drop out of its own experiment (at 8 nodes its aggregator sometimes raises "release
        # unlocked lock" and the node stops itself; the others evict it and carry on): once the
        # experiment has started everywhere, a node whose round went back to None is counted as
        # lost and the boundaries are those of the nodes still learning.
        seen = -1
        started = False
        while not stop.is_set():
            rounds = [n.state.round for n in nodes]
            if not started and all(r is not None for r in rounds):
                started = True
            live = [r for r in rounds if r is not None]
            if started and live:
                lo = min(live)
                if lo > seen:
                    seen = lo
                    for buf in flush.values():
                        buf.zero_()  # same L2 flush between rounds as the other arm
                    if lo >= W and "t0" not in marks:
                        sync_all()
                        marks["c0"] = dict(shim_pl.COUNTERS)
                        marks["t0"] = time.perf_counter()
                    if lo > W and "t0" in marks:
                        # every later boundary is a candidate end of the timed region
                        sync_all()
                        marks["node_rounds"] = marks.get("node_rounds", 0) + len(live) * (lo - W - marks.get("counted", 0))
                        marks["counted"] = lo - W
                        marks["live_last"] = len(live)
                        marks["last"] = (lo - W, time.perf_counter(), dict(shim_pl.COUNTERS))
                    if lo >= W + K and "t1" not in marks:
                        marks["t1"] = marks["last"][1]
                        return
            time.sleep(0.0005)

    sampler = bench.ClockSampler(0)
    th = threading.Thread(target=watch, daemon=True)
    nodes[0].set_start_learning(rounds=total, epochs=1)
    th.start()
    t_begin = time.time()
    started_sampler = False
    truncated = False
    while time.time() - t_begin < timeout:
        if "t0" in marks and not started_sampler:
            sampler.start()
            started_sampler = True
        if "t1" in marks:
            break
        if time.time() - t_begin > budget and "last" in marks:
            truncated = True  # out of budget: close the timed region at the last boundary seen
            break
        time.sleep(0.01)
    clocks = sampler.stop() if started_sampler else None
    stop.set()
    th.join(timeout=5)
    if truncated:
        for n in nodes:  # the reference's own API for ending an experiment early
            try:
                n.set_stop_learning()
            except Exception:
                pass
    # let the experiment end, then tear down
    end_wait = time.time() + (20 if truncated else 120)
    while any(n.state.round is not None for n in nodes) and time.time() < end_wait:
        time.sleep(0.05)
    final = {}
    try:
        final = bench.final_metrics(ref_logger.get_global_logs())
    except Exception:
        pass
    for n in nodes:
        try:
            n.stop()
        except Exception:
            pass
    if "last" not in marks:
        return {"impl": "reference", "unavailable": f"reference did not complete {W + 1} rounds in {min(timeout, budget):.0f}s"}
    done, t_end, c_end = marks["last"]
    done = min(done, K)
    ms = (t_end - marks["t0"]) * 1e3
    node_rounds = marks.get("node_rounds", n_nodes * done)  # trainer-rounds actually completed in the region
    samples = n_train * node_rounds
    value = samples / (ms / 1e3)
    h2d = (c_end["h2d"] - marks["c0"]["h2d"]) / max(node_rounds, 1)
    d2h = (c_end["d2h"] - marks["c0"]["d2h"]) / max(node_rounds, 1)
    K = done
    return {
        "metric": "fl_train_samples_per_sec",
        "value": value,
        "unit": "samples/s",
        "n_gpus": n_nodes,
        "steps": K,
        "warmup": W,
        "ms_per_step": ms / K,
        "rounds_per_sec": K / (ms / 1e3),
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": None,
        "dtype": "fp32",
        "data": "synthetic",
        "impl": "reference",
        "config": bench.bench_config(n_nodes, n_train),
        "arm": {
            "model_class": "reference mlp.py MLP (fp32, TF32 off)",
            "mode": "reference simulation mode: N nodes in one process, one GPU per node",
            "protocol": "InMemoryCommunicationProtocol",
            "settings": "pacing sleeps zeroed (--ref-fast)" if fast else "reference defaults (stock)",
            "shim": "pytorch_lightning/torchmetrics stand-ins (baseline/shim); synthetic MNIST-shaped datamodule",
            "timing": "host clock between observed round boundaries, cuda synchronize on both sides",
        },
        "truncated": truncated,
        "nodes_lost": n_nodes - int(marks.get("live_last", n_nodes)),
        "final": {"test_loss": final.get("test_loss"), "test_accuracy": final.get("test_metric"), "round": final.get("round")},
        "clocks": clocks,
        # the host-clock number above already is end to end (every batch travels host -> device, every
        # logged metric device -> host, the module moves to the device and back around each fit / test)
        "e2e": {
            "value": value,
            "unit": "samples/s",
            "ms_per_step": ms / K,
            "h2d_bytes_per_step": h2d,
            "d2h_bytes_per_step": d2h,
            "note": "per node and round, counted by the Lightning shim from the tensors it copies; the reference's own "
                    "model encode/decode copies are not counted",
        },
    }
