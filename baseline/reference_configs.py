"""The five BASELINE.json configurations on the UNMODIFIED reference (baseline/_ref).

Same conventions as ``baseline/reference_runner.py``: reference ``Node`` API, its own
in-memory protocol and simulation mode (all nodes in one process - its multi-process
gRPC mode cannot train, SURVEY Q8, and cannot carry models over 4 MiB, Q9), Lightning
shim, synthetic data.  ResNet-18 and GPT-2 do not exist in the reference; they run as
``LightningModule`` wrappers around the same ``nn.Module`` definitions this framework
uses, so the reference's learner / aggregator / gossip code is what is being measured.
"""

from __future__ import annotations

import os
import sys
import time
from typing import Any, Dict, List

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_reference_config(args: Any) -> Dict[str, Any]:
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if rank != 0:
        return {}
    ref_root = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref_root, "p2pfl")):
        return {"impl": "reference", "unavailable": "baseline/_ref is missing (run baseline/install_reference.sh)"}
    sys.path[:0] = [os.path.join(ROOT, "baseline", "shim"), ref_root]
    import pytorch_lightning as pl
    import torch
    import torch.nn.functional as F
    from p2pfl.communication.memory.memory_communication_protocol import InMemoryCommunicationProtocol
    from p2pfl.learning.pytorch.mnist_examples.models.mlp import MLP as RefMLP
    from p2pfl.management.logger import logger as ref_logger
    from p2pfl.node import Node
    from p2pfl.settings import Settings as RefSettings

    sys.path.insert(0, os.path.join(ROOT, "bench"))
    import run_config

    class LitWrapper(pl.LightningModule):
        def __init__(self, net: torch.nn.Module, amp: bool) -> None:
            super().__init__()
            self.net, self.amp = net, amp

        def forward(self, x):
            with torch.autocast(device_type=x.device.type, dtype=torch.bfloat16, enabled=self.amp and x.is_cuda):
                return self.net(x)

        def _loss(self, batch):
            x, y = batch
            out = self(x)
            if out.dim() == 3:
                out, y = out.reshape(-1, out.shape[-1]), y.reshape(-1)
            return F.cross_entropy(out.float(), y), out, y

        def configure_optimizers(self):
            return torch.optim.Adam(self.parameters(), lr=float(getattr(self.net, "lr", 1e-3)))

        def training_step(self, batch, i):
            loss, _, _ = self._loss(batch)
            self.log("train_loss", loss)
            return loss

        def validation_step(self, batch, i):
            loss, out, y = self._loss(batch)
            self.log("val_loss", loss)
            self.log("val_metric", (out.argmax(1) == y).float().mean())
            return loss

        def test_step(self, batch, i):
            loss, out, y = self._loss(batch)
            self.log("test_loss", loss)
            self.log("test_metric", (out.argmax(1) == y).float().mean())
            return loss

    n_gpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    s = run_config.spec(args.config, max(world, 1) if args.config != 1 else 1, args)
    try:
        ref_logger.set_level("WARNING")
    except Exception:
        pass
    RefSettings.TRAIN_SET_SIZE = s["train_set"]
    RefSettings.TTL = s.get("ttl", RefSettings.TTL)
    if getattr(args, "ref_fast", False):
        RefSettings.GOSSIP_MODELS_PERIOD = 0
        RefSettings.GOSSIP_PERIOD = 0
        RefSettings.WAIT_HEARTBEATS_CONVERGENCE = 0
    total = s["total"]
    nodes: List[Any] = []
    for g in range(total):
        net = s["model"](g)
        model = RefMLP() if args.config in (1, 2, 4) else LitWrapper(net, bool(s.get("amp")))
        dm = s["dm"](g, total)
        if n_gpu and s["device"] == "cuda":
            dm.device_hint = f"cuda:{(g // s['per_proc']) % n_gpu}"
        node = Node(model, dm, protocol=InMemoryCommunicationProtocol, address=f"ref-{g}")
        node.start()
        nodes.append(node)
    for a, b in run_config.edges(s["topology"], total):
        nodes[a].connect(nodes[b].addr)
    deadline = time.time() + 120
    while any(len(n.get_neighbors(only_direct=False)) != total - 1 for n in nodes) and time.time() < deadline:
        time.sleep(0.1)
    rounds = s["rounds"]
    W = min(args.warmup, max(0, rounds - 1)) if args.config != 1 else 0
    t0 = time.perf_counter()
    nodes[0].set_start_learning(rounds=rounds, epochs=s["epochs"])
    t_w = t0 if W == 0 else None
    end = time.time() + args.timeout
    started = False
    while time.time() < end:
        rs = [n.state.round for n in nodes]
        if all(r is not None for r in rs):
            started = True
            if t_w is None and min(rs) >= W:
                t_w = time.perf_counter()
        if started and all(r is None for r in rs):
            break
        time.sleep(0.002)
    if n_gpu:
        torch.cuda.synchronize()
    t1 = time.perf_counter()
    finished = all(n.state.round is None for n in nodes) and started
    per_node = s["dm"](0, total).num_samples()[0]
    for n in nodes:
        try:
            n.stop()
        except Exception:
            pass
    if not finished or t_w is None:
        return {"impl": "reference", "config": args.config, "unavailable": f"did not finish {rounds} rounds in {args.timeout}s"}
    elapsed = t1 - t_w
    timed = rounds - W
    samples = min(s["train_set"], total) * per_node * s["epochs"] * timed
    return {
        "config": args.config, "name": s["name"], "impl": "reference", "n_gpus": n_gpu, "nodes": total, "topology": s["topology"],
        "train_set": s["train_set"], "rounds_timed": timed, "warmup_rounds": W, "epochs": s["epochs"], "seconds": elapsed,
        "rounds_per_sec": timed / elapsed, "samples_per_sec": samples / elapsed, "train_samples_per_node": per_node,
        "settings": "fast" if getattr(args, "ref_fast", False) else "stock", "dtype": "bf16 autocast" if s.get("amp") else "fp32",
        "data": "synthetic", "note": "reference vote runs once (Q1): in rounds >= 1 non-trainers also train and are discarded",
    }  # fmt: skip
