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


def run_reference_experiment(n_nodes: int, steps: int, warmup: int, fast: bool, timeout: float) -> Dict[str, Any]:
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
    marks: Dict[str, float] = {}
    stop = threading.Event()

    def watch() -> None:
        # all nodes finished >= r rounds  <=>  min over nodes of state.round >= r
        while not stop.is_set():
            rounds = [n.state.round for n in nodes]
            if all(r is not None for r in rounds):
                lo = min(rounds)
                if lo >= W and "t0" not in marks:
                    if n_gpu:
                        torch.cuda.synchronize()
                    marks["t0"] = time.perf_counter()
                if lo >= W + K and "t1" not in marks:
                    if n_gpu:
                        torch.cuda.synchronize()
                    marks["t1"] = time.perf_counter()
                    return
            time.sleep(0.0005)

    sampler = bench.ClockSampler(0)
    th = threading.Thread(target=watch, daemon=True)
    nodes[0].set_start_learning(rounds=total, epochs=1)
    th.start()
    t_begin = time.time()
    started_sampler = False
    while time.time() - t_begin < timeout:
        if "t0" in marks and not started_sampler:
            sampler.start()
            started_sampler = True
        if "t1" in marks:
            break
        time.sleep(0.01)
    clocks = sampler.stop() if started_sampler else None
    stop.set()
    # let the experiment end, then tear down
    end_wait = time.time() + 120
    while any(n.state.round is not None for n in nodes) and time.time() < end_wait:
        time.sleep(0.05)
    for n in nodes:
        try:
            n.stop()
        except Exception:
            pass
    if "t1" not in marks:
        return {"impl": "reference", "unavailable": f"reference did not complete {W + K} rounds in {timeout}s"}
    ms = (marks["t1"] - marks["t0"]) * 1e3
    samples = n_nodes * n_train * K
    return {
        "metric": "fl_train_samples_per_sec",
        "value": samples / (ms / 1e3),
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
        "config": {
            "model": "MNIST-shaped MLP 784-256-128-10 (reference mlp.py)",
            "nodes": n_nodes,
            "topology": "full",
            "aggregator": "FedAvg, all nodes in the train set",
            "epochs_per_round": 1,
            "train_samples_per_node": n_train,
            "global_batch": bench.BATCH * n_nodes,
            "seq_len": None,
            "parallelism": f"fl{n_nodes} (reference simulation mode: N nodes in one process, one GPU per node)",
            "protocol": "InMemoryCommunicationProtocol",
            "settings": "pacing sleeps zeroed (--ref-fast)" if fast else "reference defaults (stock)",
            "shim": "pytorch_lightning/torchmetrics stand-ins (baseline/shim); synthetic MNIST-shaped datamodule",
            "timing": "host clock between observed round boundaries, cuda synchronize on both sides",
        },
        "clocks": clocks,
    }
