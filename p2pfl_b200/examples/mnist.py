"""N nodes on a line learn an MNIST-shaped MLP with FedAvg (reference: p2pfl/examples/mnist.py)."""

from __future__ import annotations

import argparse
import json
import os
import sys
import time
import uuid

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from p2pfl_b200.learning.data import MnistFederatedDM  # noqa: E402
from p2pfl_b200.management.logger import logger  # noqa: E402
from p2pfl_b200.models import CNN, MLP  # noqa: E402
from p2pfl_b200.node import Node  # noqa: E402
from p2pfl_b200.utils import check_equal_models, wait_4_results, wait_convergence  # noqa: E402


def parse_args() -> argparse.Namespace:
    p = argparse.ArgumentParser(description="p2pfl_b200 MNIST-shaped experiment")
    p.add_argument("--nodes", type=int, default=2)
    p.add_argument("--rounds", type=int, default=2)
    p.add_argument("--epochs", type=int, default=0)
    p.add_argument("--model", choices=["mlp", "cnn"], default="mlp")
    p.add_argument("--topology", choices=["line", "ring", "full", "star"], default="line")
    p.add_argument("--show_metrics", action="store_true", help="write the collected metrics as JSON (no plot windows)")
    p.add_argument("--metrics_file", default="metrics.json")
    p.add_argument("--measure_time", action="store_true")
    p.add_argument("--use_unix_socket", action="store_true")
    p.add_argument("--use_local_protocol", action="store_true", help="in-process transport instead of gRPC")
    p.add_argument("--device_plane", action="store_true", help="NVLink data plane + fused learner (needs GPUs)")
    p.add_argument("--token", default="", help="API token of the web dashboard")
    p.add_argument("--web_url", default="")
    p.add_argument("--samples", type=int, default=60000, help="size of the synthetic sample pool")
    return p.parse_args()


def build_nodes(args: argparse.Namespace) -> list:
    import torch

    n_gpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    learner = None
    if args.device_plane:
        from p2pfl_b200.communication.nvlink import NvlinkMemoryProtocol as protocol
        from p2pfl_b200.learning.b200_learner import B200Learner as learner
    elif args.use_local_protocol:
        from p2pfl_b200.communication.memory import InMemoryCommunicationProtocol as protocol
    else:
        from p2pfl_b200.communication.grpc import GrpcCommunicationProtocol as protocol
    model_cls = MLP if args.model == "mlp" else CNN
    nodes = []
    for i in range(args.nodes):
        if args.use_unix_socket and not (args.use_local_protocol or args.device_plane):
            address = f"unix:///tmp/p2pfl-{uuid.uuid4().hex[:12]}.sock"
        elif args.use_local_protocol or args.device_plane:
            address = f"node-{i}"
        else:
            address = "127.0.0.1"
        kwargs = {"device": f"cuda:{i % n_gpu}"} if n_gpu else {"device": "cpu"}
        node = Node(
            model_cls(),
            MnistFederatedDM(i, args.nodes, n_train=args.samples, n_test=args.samples // 6),
            address=address,
            protocol=protocol,
            learner=learner,
            **kwargs,
        )
        node.start()
        nodes.append(node)
    return nodes


def connect(nodes: list, topology: str) -> None:
    n = len(nodes)
    if topology == "full":
        pairs = [(i, j) for i in range(n) for j in range(i)]
    elif topology == "star":
        pairs = [(i, 0) for i in range(1, n)]
    else:
        pairs = [(i + 1, i) for i in range(n - 1)]
        if topology == "ring" and n > 2:
            pairs.append((0, n - 1))
    for a, b in pairs:
        nodes[a].connect(nodes[b].addr)
        time.sleep(0.05)
    wait_convergence(nodes, n - 1, only_direct=False, wait=30)


def main() -> None:
    args = parse_args()
    if args.token and args.web_url:
        logger.connect_web(args.web_url, args.token)
    t0 = time.time()
    nodes = build_nodes(args)
    try:
        connect(nodes, args.topology)
        nodes[0].set_start_learning(rounds=args.rounds, epochs=args.epochs)
        wait_4_results(nodes)
        check_equal_models(nodes, atol=1e-1)
        if args.show_metrics:
            with open(args.metrics_file, "w") as f:
                json.dump({"local": logger.get_local_logs(), "global": logger.get_global_logs()}, f, indent=1, default=str)
            print(f"metrics written to {args.metrics_file}")
            for exp, per_node in logger.get_global_logs().items():
                for node, metrics in per_node.items():
                    for name, series in metrics.items():
                        print(f"{exp} {node} {name}: " + " ".join(f"r{r}={v:.4f}" for r, v in series))
    finally:
        for node in nodes:
            node.stop()
    if args.measure_time:
        print(f"--- {time.time() - t0:.3f} seconds ---")


if __name__ == "__main__":
    main()
