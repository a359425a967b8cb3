#!/usr/bin/env python
"""Drivers for the five BASELINE.json configurations.

    python bench/run_config.py --config 1                       # 2 nodes, in-memory, CPU
    torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 bench/run_config.py --config 2|3|4|5
    ... --impl reference                                        # the same config on baseline/_ref

  1  MNIST-shaped MLP, 2 nodes, in-memory protocol on CPU, FedAvg, 1 round x 1 epoch
  2  MNIST-shaped MLP, one node per GPU, fully connected, FedAvg, 10 rounds
  3  CIFAR-10-shaped ResNet-18, one node per GPU, ring, partial aggregates, bf16
  4  64 simulated nodes (8 per GPU), MNIST-shaped MLP, line, vote selects 16 trainers per round
  5  GPT-2 small (124M), one node per GPU, full-model gossip + FedAvg (exchange stress)

Prints one JSON line: rounds/s, samples/s (max-over-ranks time between round
boundaries, cuda-synchronised), and the data-plane byte counters.
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import time
from typing import Any, Dict, List

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def env() -> Dict[str, int]:
    return {"rank": int(os.environ.get("RANK", 0)), "local": int(os.environ.get("LOCAL_RANK", 0)), "world": int(os.environ.get("WORLD_SIZE", 1))}


def spec(config: int, world: int, args: argparse.Namespace) -> Dict[str, Any]:
    from p2pfl_b200.learning.data import Cifar10FederatedDM, MnistFederatedDM, TokenFederatedDM
    from p2pfl_b200.models import MLP, ResNet18, gpt2_small

    if config == 1:
        return dict(total=2, per_proc=2, model=lambda i: MLP(seed=i), dm=lambda i, n: MnistFederatedDM(i, n, n_train=6000 * n, n_test=1000 * n),
                    topology="line", train_set=2, rounds=1, epochs=1, device="cpu", plane="memory", name="mlp 2 nodes in-memory CPU")  # fmt: skip
    if config == 2:
        return dict(total=world, per_proc=1, model=lambda i: MLP(seed=i), dm=lambda i, n: MnistFederatedDM(i % 8, 8),
                    topology="full", train_set=world, rounds=args.rounds or 10, epochs=1, device="cuda", plane="nvlink", name="mlp full fedavg")  # fmt: skip
    if config == 3:
        return dict(total=world, per_proc=1, model=lambda i: ResNet18(seed=i), dm=lambda i, n: Cifar10FederatedDM(i % 8, 8, n_train=args.samples or 8192, n_test=2048),
                    topology="ring", train_set=min(4, world), rounds=args.rounds or 3, epochs=1, device="cuda", plane="nvlink", amp=True, name="resnet18 ring partial bf16")  # fmt: skip
    if config == 4:
        per = args.per_gpu or 8
        return dict(total=world * per, per_proc=per, model=lambda i: MLP(seed=i), dm=lambda i, n: MnistFederatedDM(i % 64, 64),
                    topology="line", train_set=min(16, world * per), rounds=args.rounds or 3, epochs=1, device="cuda", plane="nvlink", ttl=max(10, world * per),
                    name="mlp 64 simulated nodes line vote-16")  # fmt: skip
    if config == 5:
        return dict(total=world, per_proc=1, model=lambda i: gpt2_small(seed=i), dm=lambda i, n: TokenFederatedDM(i, n, batch_size=4, seq_len=1024, n_train=(args.samples or 16) * n, n_test=4 * n),
                    topology="full", train_set=world, rounds=args.rounds or 3, epochs=args.epochs if args.epochs is not None else 1, device="cuda", plane="nvlink", amp=True,
                    name="gpt2-small full-model gossip fedavg")  # fmt: skip
    raise SystemExit(f"unknown config {config}")


def edges(topology: str, n: int) -> List[tuple]:
    if topology == "full":
        return [(i, j) for i in range(n) for j in range(i)]
    e = [(i + 1, i) for i in range(n - 1)]
    if topology == "ring" and n > 2:
        e.append((0, n - 1))
    return e


def run_ours(args: argparse.Namespace) -> Dict[str, Any]:
    e = env()
    rank, local, world = e["rank"], e["local"], e["world"]
    s = spec(args.config, world, args)
    on_gpu = s["device"] == "cuda"
    if on_gpu:
        torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl" if on_gpu else "gloo", **({"device_id": torch.device(f"cuda:{local}")} if on_gpu else {}))

    from p2pfl_b200 import ops as _ops
    from p2pfl_b200.management.logger import logger

    if on_gpu:
        _ops.set_home_device(local)
    _ops.thread_uses_cuda()
    from p2pfl_b200.node import Node
    from p2pfl_b200.settings import Settings
    from p2pfl_b200.utils import wait_convergence

    logger.set_level("WARNING")
    Settings.LOG_LEVEL = "WARNING"
    Settings.TRAIN_SET_SIZE = s["train_set"]
    Settings.DEVICE_INBOX_SLOTS = max(2, min(16, s["train_set"]))
    Settings.TTL = s.get("ttl", Settings.TTL)
    Settings.GOSSIP_MODELS_PERIOD = 0.05
    Settings.GOSSIP_PERIOD = 0.0
    Settings.GOSSIP_MODELS_PER_ROUND = 4
    Settings.WAIT_HEARTBEATS_CONVERGENCE = 0.3
    Settings.HEARTBEAT_PERIOD = 2.0
    Settings.HEARTBEAT_TIMEOUT = 60.0
    Settings.GRPC_TIMEOUT = 60.0
    Settings.AGGREGATION_TIMEOUT = 300.0
    Settings.VOTE_TIMEOUT = 300.0
    Settings.AMP_BF16 = bool(s.get("amp", False))
    Settings.GRPC_SERVER_WORKERS = 8

    total, per = s["total"], s["per_proc"]
    port0 = int(os.environ.get("MASTER_PORT", "29500")) + 300
    if s["plane"] == "memory":
        from p2pfl_b200.communication.memory import InMemoryCommunicationProtocol as Proto
        from p2pfl_b200.learning.torch_learner import TorchLearner as Learner

        addr = lambda g: f"cfg-node-{g}"  # noqa: E731
    else:
        if args.control == "grpc":
            from p2pfl_b200.communication.nvlink import NvlinkCommunicationProtocol as Proto
        else:
            from p2pfl_b200.communication.nvlink import NvlinkSocketProtocol as Proto
        from p2pfl_b200.learning.b200_learner import B200Learner as Learner

        addr = lambda g: f"127.0.0.1:{port0 + g}"  # noqa: E731
    mine = list(range(rank * per, (rank + 1) * per)) if world > 1 else list(range(total))
    nodes = {}
    for g in mine:
        kw = {"device": f"cuda:{local}"} if on_gpu else {"device": "cpu"}
        nodes[g] = Node(s["model"](g), s["dm"](g, total), address=addr(g), learner=Learner, protocol=Proto, **kw)
        nodes[g].start()
    if world > 1:
        dist.barrier()
    for a, b in edges(s["topology"], total):
        if a in nodes:
            nodes[a].connect(addr(b))
    if world > 1:
        dist.barrier()
    wait_convergence(list(nodes.values()), total - 1, wait=120, only_direct=False)
    if world > 1:
        dist.barrier()

    rounds, W = s["rounds"], min(args.warmup, max(0, s["rounds"] - 1)) if args.config != 1 else 0
    marks: Dict[str, float] = {}
    n_local = len(nodes)
    done_rounds: Dict[int, int] = {}

    def listener(r: int) -> None:
        done_rounds[r] = done_rounds.get(r, 0) + 1
        if done_rounds[r] == n_local and r == W and "t0" not in marks:
            # (no device-wide synchronize here: another node of this process may be capturing its
            # epoch graph, and CUDA refuses a device synchronize while any stream captures)
            marks["t0"] = time.perf_counter()

    for n in nodes.values():
        n.state.round_listeners = [listener]
    if W == 0:
        marks["t0"] = time.perf_counter()
    t_start = time.perf_counter()
    if 0 in nodes:
        nodes[0].set_start_learning(rounds=rounds, epochs=s["epochs"])
    for n in nodes.values():
        while n.state.finished.is_set() and time.perf_counter() - t_start < 60:
            time.sleep(0.001)
    for n in nodes.values():
        if not n.wait_learning_finished(args.timeout):
            raise SystemExit(f"rank {rank}: node {n.addr} did not finish")
    if on_gpu:
        torch.cuda.synchronize()
    t1 = time.perf_counter()
    stats = {"fused": 0, "pulled": 0, "host": 0, "fused_bytes": 0, "pulled_bytes": 0}
    train_samples = 0
    trained_rounds = 0
    for n in nodes.values():
        for k in stats:
            stats[k] += getattr(n._communication_protocol, "stats", {}).get(k, 0)
        lr = n.state.learner
        train_samples += lr.get_num_samples()[0]
        trained_rounds += n.learning_workflow.history.count("TrainStage")
    elapsed = t1 - marks["t0"]
    vec = torch.tensor([elapsed, float(trained_rounds), float(train_samples)] + [float(stats[k]) for k in sorted(stats)], dtype=torch.float64)
    if world > 1:
        vec = vec.cuda() if on_gpu else vec
        mx = vec.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = vec.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        elapsed = float(mx[0])
        trained_rounds = int(sm[1])
        stats = {k: int(sm[3 + i]) for i, k in enumerate(sorted(stats))}
    for n in nodes.values():
        n.stop()
    timed_rounds = rounds - W
    per_node_samples = next(iter(nodes.values())).state.learner.get_num_samples()[0] if nodes else 0
    # samples actually trained in the timed rounds = (trainer-rounds in timed window) x shard x epochs
    trainers_per_round = min(s["train_set"], total)
    samples = trainers_per_round * per_node_samples * s["epochs"] * timed_rounds
    out = {
        "config": args.config, "name": s["name"], "impl": "p2pfl_b200", "n_gpus": world if on_gpu else 0, "nodes": total,
        "topology": s["topology"], "train_set": s["train_set"], "rounds_timed": timed_rounds, "warmup_rounds": W, "epochs": s["epochs"],
        "seconds": elapsed, "rounds_per_sec": timed_rounds / elapsed, "samples_per_sec": samples / elapsed,
        "train_samples_per_node": per_node_samples, "data_plane": stats,
        "gossip_GB_total": (stats["fused_bytes"] + stats["pulled_bytes"]) / 1e9,
        "dtype": "bf16" if s.get("amp") or args.config in (2, 4) else "fp32", "data": "synthetic",
    }  # fmt: skip
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out if rank == 0 else {}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, required=True)
    ap.add_argument("--rounds", type=int, default=0)
    ap.add_argument("--epochs", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--samples", type=int, default=0)
    ap.add_argument("--per-gpu", type=int, default=0)
    ap.add_argument("--timeout", type=float, default=900.0)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ref-fast", action="store_true")
    ap.add_argument("--control", default="socket", choices=["socket", "grpc"])
    args = ap.parse_args()
    if args.impl == "reference":
        from baseline.reference_configs import run_reference_config

        res = run_reference_config(args)
    else:
        res = run_ours(args)
    if env()["rank"] == 0:
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
