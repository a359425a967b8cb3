"""Two-process quickstart, process 1: start a node and wait for a peer (reference: p2pfl/examples/node1.py)."""

from __future__ import annotations

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from p2pfl_b200.learning.data import MnistFederatedDM  # noqa: E402
from p2pfl_b200.models import MLP  # noqa: E402
from p2pfl_b200.node import Node  # noqa: E402


def main() -> None:
    p = argparse.ArgumentParser()
    p.add_argument("--port", type=int, default=6666)
    args = p.parse_args()
    node = Node(MLP(), MnistFederatedDM(0, 2), address=f"127.0.0.1:{args.port}")
    node.start()
    input("Press ENTER to stop the node\n")
    node.stop()


if __name__ == "__main__":
    main()
