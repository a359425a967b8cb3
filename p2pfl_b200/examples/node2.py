"""Two-process quickstart, process 2: connect to node1 and start learning (reference: p2pfl/examples/node2.py)."""

from __future__ import annotations

import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from p2pfl_b200.learning.data import MnistFederatedDM  # noqa: E402
from p2pfl_b200.models import MLP  # noqa: E402
from p2pfl_b200.node import Node  # noqa: E402


def main() -> None:
    p = argparse.ArgumentParser()
    p.add_argument("--port", type=int, default=6666, help="port of node1")
    p.add_argument("--rounds", type=int, default=2)
    p.add_argument("--epochs", type=int, default=1)
    args = p.parse_args()
    node = Node(MLP(), MnistFederatedDM(1, 2), address="127.0.0.1")
    node.start()
    node.connect(f"127.0.0.1:{args.port}")
    time.sleep(4)
    node.set_start_learning(rounds=args.rounds, epochs=args.epochs)
    while not node.wait_learning_finished(1.0):
        pass
    node.stop()


if __name__ == "__main__":
    main()
