"""Debug aid: co-located trainers plus a waiting node on one GPU; prints a checksum per node and round."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from p2pfl_b200.communication.nvlink import NvlinkMemoryProtocol
from p2pfl_b200.learning.b200_learner import B200Learner
from p2pfl_b200.learning.data import MnistFederatedDM
from p2pfl_b200.models import MLP
from p2pfl_b200.node import Node
from p2pfl_b200.settings import Settings, set_test_settings
from p2pfl_b200.utils import wait_4_results, wait_convergence

set_test_settings()
Settings.LOG_LEVEL = "WARNING"
from p2pfl_b200.management.logger import logger
logger.set_level("WARNING")
n, ts, rounds = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
Settings.TRAIN_SET_SIZE = ts
nodes = [Node(MLP(seed=i), MnistFederatedDM(i, n, n_train=640 * n, n_test=128 * n), protocol=NvlinkMemoryProtocol,
              learner=B200Learner, device="cuda:0") for i in range(n)]
for x in nodes:
    x.start()
for i in range(1, n):
    for j in range(i):
        nodes[i].connect(nodes[j].addr)
wait_convergence(nodes, n - 1, only_direct=True)
log = []
def mk(node):
    def fn(r):
        lr = node.state.learner
        lr.stream.synchronize(); torch.cuda.synchronize()
        log.append((time.time(), node.addr, r, float(lr.flat.double().sum()), sorted(node.state.train_set)))
    return fn
for x in nodes:
    x.state.round_listeners = [mk(x)]
nodes[0].set_start_learning(rounds=rounds, epochs=1)
wait_4_results(nodes, 120)
for t, a, r, c, tsx in sorted(log, key=lambda e: (e[2], e[1])):
    print(r, a, f"{c:.9f}", tsx)
for x in nodes:
    print(x.addr, dict(x._communication_protocol.stats))
    x.stop()
