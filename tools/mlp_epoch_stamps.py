"""Bring-up aid (needs `python build.py --bringup`): per-phase SM-clock stamps of one steady-state
step of the whole-epoch MLP kernel, per CTA, and the CUDA-event time of the epoch."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from p2pfl_b200 import ops
from p2pfl_b200.learning.b200_learner import B200Learner
from p2pfl_b200.learning.data import MnistFederatedDM
from p2pfl_b200.models import MLP

dm = MnistFederatedDM(0, 8, n_train=60000, n_test=10000)
lr = B200Learner(MLP(seed=0), dm, "prof", epochs=1, device="cuda:0")
lr.fit(); lr.fit()
stamps = torch.zeros(16, 32, dtype=torch.int64, device="cuda:0")
rc = ops.lib().p2pfl_mlp_epoch_set_stamps(ctypes.c_void_p(stamps.data_ptr()))
ep = next(v for k, v in lr._graphs.items() if k[0] == "epoch")
n = ep["n"]
with torch.cuda.stream(lr.stream):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        e0.record()
        ops.mlp_epoch(ep["xs"], ep["ys"], n, lr.flat, lr.adam_m, lr.adam_v, lr.w16, ep["loss"], lr.adam_step, 1e-3)
        e1.record()
lr.stream.synchronize()
ms = e0.elapsed_time(e1)
print(f"epoch kernel: {ms:.3f} ms for {ep['nb']} steps = {ms * 1e3 / ep['nb']:.2f} us/step (stamps rc={rc})")
st = stamps.cpu()
names = ["start", "fwd1 done", "H1 stored", "barrier a", "fwd2 end", "fwd3+CE end", "dgrad3 end", "dgrad2 end", "wgrad done", "adam done", "fences", "barrier b"]
hz = ops.rt_sm_clock_hz(0)
fine = ["issuer awake", "issued+commit", "warp0 woke", "tmem ld", "math+st", "fence"]
for c in (0, 8):
    t0 = int(st[c, 5])
    print(f"CTA {c} dgrad3 link (us after the previous link's end): " + "  ".join(f"{fine[i]}={(int(st[c, 16 + i]) - t0) / hz * 1e6:.2f}" for i in range(6)) + f"  link end={(int(st[c, 6]) - t0) / hz * 1e6:.2f}")
for c in (0, 7, 8, 15):
    t0 = int(st[c, 0])
    print(f"CTA {c}: " + "  ".join(f"{names[i]}={(int(st[c, i]) - t0) / hz * 1e6:.2f}" for i in range(1, 12) if int(st[c, i])))
