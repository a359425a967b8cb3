#!/usr/bin/env python
"""Small single-GPU workload for ncu: a few fused train steps (graph replay), an
evaluation, a FedAvg, and a GPT-2-sized push + reduce on one device.  Run it plain
first; only then under ncu (B200_PROFILING.md)."""

import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from p2pfl_b200 import ops  # noqa: E402
from p2pfl_b200.learning.b200_learner import B200Learner  # noqa: E402
from p2pfl_b200.learning.data import MnistFederatedDM  # noqa: E402
from p2pfl_b200.models import MLP  # noqa: E402
from p2pfl_b200.parallel.arena import DeviceArena  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "all"
torch.cuda.set_device(0)
if what in ("all", "step"):
    dm = MnistFederatedDM(0, 1, n_train=1280, n_test=256)
    learner = B200Learner(MLP(seed=0), dm, "prof", epochs=1, device="cuda:0")
    learner.fit()  # builds + warms the graphs
    learner.fit()
    print("eval", learner.evaluate())
    torch.cuda.synchronize()
if what == "mlpstep":
    # the fused MLP step in every launch arrangement (per-link, single-CTA chain, cluster chains)
    from p2pfl_b200.settings import Settings  # noqa: E402

    dm = MnistFederatedDM(0, 1, n_train=640, n_test=128)
    x, y = dm.tensors("train")
    for chain, cluster in ((False, 0), (True, 0), (False, 2), (False, 4), (False, 8), (False, 16)):
        Settings.MLP_CHAIN, Settings.MLP_CHAIN_CLUSTER = chain, cluster
        learner = B200Learner(MLP(seed=0), dm, "prof", epochs=1, device="cuda:0")
        stamps = torch.zeros(24, 8, dtype=torch.int64, device="cuda:0")
        ops.lib().p2pfl_gemm_chain_set_stamps(ctypes.c_void_p(stamps.data_ptr()))
        for i in range(3):
            learner.debug_single_step(x[i * 32 : (i + 1) * 32], y[i * 32 : (i + 1) * 32])
        torch.cuda.synchronize()
        ops.lib().p2pfl_gemm_chain_set_stamps(ctypes.c_void_p(0))
        if chain or cluster:
            st = stamps.cpu()
            t0 = int(st[0, 0])
            print(f"chain={chain} cluster={cluster}: per-link stamps in us (1.9 GHz) relative to link 0 start")
            for p_ in range(8):
                if int(st[p_, 0]) == 0:
                    continue
                print("  link", p_, [round((int(v) - t0) / 1900.0, 2) if int(v) else None for v in st[p_, :7]],
                      "epilogue", [round((int(v) - t0) / 1900.0, 2) if int(v) else None for v in st[12 + p_, :6]])
        learner.close()
    print("mlpstep ok")
if what == "resnet":
    # the graph build runs every kernel of the step once outside capture, so a launch list
    # under ncu shows the whole fused ResNet-18 step
    from p2pfl_b200.learning.data import Cifar10FederatedDM  # noqa: E402
    from p2pfl_b200.models import ResNet18  # noqa: E402

    dm = Cifar10FederatedDM(0, 1, n_train=640, n_test=64, batch_size=64)
    learner = B200Learner(ResNet18(seed=0), dm, "prof", epochs=1, device="cuda:0")
    learner.fit()
    ev0 = torch.cuda.Event(enable_timing=True)
    ev1 = torch.cuda.Event(enable_timing=True)
    ev0.record(learner.stream)
    for _ in range(5):
        learner.fit()
    ev1.record(learner.stream)
    torch.cuda.synchronize()
    print("resnet ms/fit", ev0.elapsed_time(ev1) / 5, "steps/fit", learner._round_steps,
          "launches/epoch", learner.kernel_launches_per_epoch)
if what in ("all", "gossip"):
    n, tile = 124_441_344, 16384
    a = DeviceArena(n, slots=2, tile=tile, wire_bf16=True, device=0)
    b = DeviceArena(n, slots=2, tile=tile, wire_bf16=True, device=0)
    a.params.normal_()
    b.params.normal_()
    w16 = torch.empty(n, dtype=torch.bfloat16, device="cuda:0")
    for seq in (1, 3, 5):
        push = ops.make_push_table([b.slot_ptr(0, 1)], [b.flag_ptr(0, 1)], True)
        ops.gossip_push(a.params, tile, seq, push)
        red = ops.make_reduce_table([b.slot_ptr(0, 1)], [b.flag_ptr(0, 1)], [0.5], 1, 0.5, True, True)
        ops.gossip_reduce(b.params, w16, tile, seq, red, None, 0, None)
    torch.cuda.synchronize()
    print("gossip ok")
if what in ("all", "gemm"):
    M, N, K = 4096, 4096, 4096
    A = (torch.randn(M, K, device="cuda:0") / 64).to(torch.bfloat16)
    B = torch.randn(N, K, device="cuda:0").to(torch.bfloat16)
    out = torch.empty(M, N, device="cuda:0", dtype=torch.bfloat16)
    for _ in range(3):
        ops.gemm_tn(A, K, B, K, M, N, K, out_bf16=out, ld_bf16=N)
    torch.cuda.synchronize()
    print("gemm ok")

if what == "evidence":
    # One pass over every kernel family for the per-kernel ncu summaries (profiles/r2/ncu_*.md):
    # the whole-epoch MLP kernel, the CNN and ResNet steps (GEMM, im2col, BatchNorm, pooling, Adam),
    # evaluation, FedAvg, the co-located exchange (one cooperative launch) and the copy-engine push.
    from p2pfl_b200.learning.data import Cifar10FederatedDM  # noqa: E402
    from p2pfl_b200.models import CNN, ResNet18  # noqa: E402
    from p2pfl_b200.parallel.arena import emulate_exchange  # noqa: E402

    dm = MnistFederatedDM(0, 1, n_train=1280, n_test=256)
    lr = B200Learner(MLP(seed=0), dm, "prof", epochs=1, device="cuda:0")
    lr.fit()
    lr.fit()
    print("mlp eval", lr.evaluate())
    cnn = B200Learner(CNN(seed=0), MnistFederatedDM(0, 1, n_train=256, n_test=64), "prof-cnn", epochs=1, device="cuda:0")
    cnn.fit()
    print("cnn eval", cnn.evaluate())
    rn = B200Learner(ResNet18(seed=0), Cifar10FederatedDM(0, 1, n_train=128, n_test=64, batch_size=64), "prof-rn", epochs=1, device="cuda:0")
    rn.fit()
    print("resnet eval", rn.evaluate())
    n = 11_173_984 // 32 * 32
    models = [torch.randn(n, device="cuda:0") for _ in range(4)]
    out = torch.empty(n, device="cuda:0")
    ops.fedavg_into(models, [0.25] * 4, out, None)
    emulate_exchange(models, [1, 2, 3, 4])
    a = DeviceArena(n, slots=2, tile=16384, wire_bf16=True, device=0)
    b = DeviceArena(n, slots=2, tile=16384, wire_bf16=True, device=0)
    a.params.copy_(models[0])
    ops.gossip_push_bulk(a.params, 3, ops.make_push_table([b.slot_ptr(0, 1)], [b.flag_ptr(0, 1)], True))
    torch.cuda.synchronize()
    print("evidence ok")
