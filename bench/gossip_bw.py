#!/usr/bin/env python
"""Gossip data-plane microbenchmark: fused push+FedAvg kernel vs NCCL all-gather+average.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 bench/gossip_bw.py [--model gpt2]

Every rank owns a device arena, maps the others through CUDA IPC and runs the
all-to-all exchange of a model of the named size (BASELINE.json config 5 for
``gpt2``).  Times are CUDA events on the launching stream, max over ranks, after
warm-up; inputs exceed L2 for the large models, the small ones get an L2 flush.
Roofline = bytes that must enter each GPU over NVLink / 770 GB/s measured peer
bandwidth (B200_PROFILING.md).  Prints one JSON line per (model, wire dtype).
"""

from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SIZES = {"mlp": 235_168, "resnet18": 11_173_984, "gpt2": 124_441_344}
NVLINK_GBS = 770.0


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--models", default="mlp,resnet18,gpt2")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--tile", type=int, default=16384)
    args = ap.parse_args()
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    from p2pfl_b200 import ops
    from p2pfl_b200.parallel.arena import DeviceArena, PeerArena

    dev = torch.device(f"cuda:{local}")
    flush = torch.empty(64 << 20, dtype=torch.float32, device=dev)
    for name in args.models.split(","):
        n = (SIZES[name] + 31) // 32 * 32
        for wire_bf16 in (True, False):
            arena = DeviceArena(n, slots=world, tile=args.tile, wire_bf16=wire_bf16, device=local)
            handles = [None] * world
            dist.all_gather_object(handles, (arena.ipc_handle(), local))
            peers = {r: PeerArena.open_ipc(h, arena.geo, local, d) for r, (h, d) in enumerate(handles) if r != rank}
            others = [r for r in range(world) if r != rank]
            torch.manual_seed(rank)
            model = torch.randn(n, device=dev)
            w16 = torch.empty(n, dtype=torch.bfloat16, device=dev)
            weights = [1.0 / world] * world
            status = torch.zeros(4, dtype=torch.int32, device=dev)

            def fused(seq: int) -> None:
                bank = seq & 1
                push = ops.make_push_table([peers[r].slot_ptr(rank, bank) for r in others],
                                           [peers[r].flag_ptr(rank, bank) for r in others], wire_bf16)
                red = ops.make_reduce_table([arena.slot_ptr(r, bank) for r in others],
                                            [arena.flag_ptr(r, bank) for r in others],
                                            [weights[r] for r in others], rank, weights[rank], wire_bf16, True)
                ops.gossip_fused(arena.params, w16, args.tile, seq, push, red, None, int(20e9), status)

            # correctness once: every rank must hold the same average
            arena.params.copy_(model)
            dist.barrier()
            fused(1)
            torch.cuda.synchronize()
            ref = model.to(torch.bfloat16).float() if wire_bf16 else model.clone()
            gathered = [torch.empty_like(ref) for _ in range(world)]
            dist.all_gather(gathered, ref)
            expect = torch.zeros_like(ref)
            for g in gathered:
                expect += g / world
            err = (arena.params - expect).abs().max().item()

            times = []
            seq = 2
            for it in range(args.warmup + args.iters):
                arena.params.copy_(model)
                flush.zero_()
                torch.cuda.synchronize()
                dist.barrier()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fused(seq)
                e1.record()
                torch.cuda.synchronize()
                seq += 1
                if it >= args.warmup:
                    times.append(e0.elapsed_time(e1))
            t = torch.tensor([sum(times) / len(times), min(times)], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)

            # baseline: NCCL all-gather (same wire dtype) + average
            wire = model.to(torch.bfloat16) if wire_bf16 else model
            buf = torch.empty(world * n, dtype=wire.dtype, device=dev)
            out = torch.empty(n, device=dev)
            btimes = []
            for it in range(args.warmup + args.iters):
                flush.zero_()
                torch.cuda.synchronize()
                dist.barrier()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                dist.all_gather_into_tensor(buf, wire)
                ops.fedavg_into([buf[r * n : (r + 1) * n] for r in range(world)], weights, out, w16)
                e1.record()
                torch.cuda.synchronize()
                if it >= args.warmup:
                    btimes.append(e0.elapsed_time(e1))
            bt = torch.tensor([sum(btimes) / len(btimes), min(btimes)], device=dev, dtype=torch.float64)
            dist.all_reduce(bt, op=dist.ReduceOp.MAX)

            wb = 2 if wire_bf16 else 4
            bytes_in = (world - 1) * n * wb
            roof_ms = bytes_in / (NVLINK_GBS * 1e9) * 1e3
            if rank == 0:
                print(json.dumps({
                    "bench": "gossip_fused", "model": name, "elements": n, "n_gpus": world,
                    "wire": "bf16" if wire_bf16 else "fp32", "max_abs_err_vs_allgather_avg": err,
                    "fused_ms_mean": float(t[0]), "fused_ms_best": float(t[1]),
                    "gossip_GBps_in_per_gpu": bytes_in / (float(t[0]) * 1e-3) / 1e9,
                    "roofline_ms_at_770GBps": roof_ms, "fraction_of_nvlink_roofline": roof_ms / float(t[0]),
                    "nccl_allgather_avg_ms_mean": float(bt[0]), "nccl_allgather_avg_ms_best": float(bt[1]),
                    "speedup_vs_nccl": float(bt[0]) / float(t[0]),
                }), flush=True)
            for p in peers.values():
                p.close()
            dist.barrier()
            arena.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
