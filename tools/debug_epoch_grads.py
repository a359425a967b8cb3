"""Debug aid: one step through the epoch kernel vs the per-link path; where do the gradients differ?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from p2pfl_b200.learning.b200_learner import B200Learner
from p2pfl_b200.learning.data import MnistFederatedDM
from p2pfl_b200.models import MLP
from p2pfl_b200.settings import Settings

out = {}
for ek in (True, False):
    Settings.MLP_EPOCH_KERNEL, Settings.MLP_CHAIN_CLUSTER = ek, 0
    dm = MnistFederatedDM(0, 1, n_train=640, n_test=128)
    lr = B200Learner(MLP(seed=3), dm, "p", epochs=1, device="cuda:0")
    x, y = dm.tensors("train")
    lr.debug_single_step(x[:32], y[:32])
    out[ek] = {k: v.clone() for k, v in lr.layout.views(lr.grads).items()}
    lr.close()
for name in out[True]:
    a, b = out[True][name], out[False][name]
    print(name, tuple(a.shape), "rel", float((a - b).norm() / (b.norm() + 1e-12)))
a, b = out[True]["l2.weight"], out[False]["l2.weight"]
err = (a - b).abs()
print("row blocks of 8 (max err):", [round(float(err[8 * i : 8 * i + 8].max()), 5) for i in range(16)])
print("col blocks of 32 (max err):", [round(float(err[:, 32 * i : 32 * i + 32].max()), 5) for i in range(8)])
print("ref max", float(b.abs().max()))
# is a row of `a` equal to another row of `b`?
for r in (0, 8, 9, 17):
    d = ((b - a[r : r + 1]) ** 2).sum(1)
    print("row", r, "of epoch-kernel grad is closest to reference row", int(d.argmin()), float(d.min()))
