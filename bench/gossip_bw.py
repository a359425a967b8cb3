#!/usr/bin/env python
"""Gossip data-plane microbenchmark: fused push+FedAvg kernel vs NCCL all-gather+average.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 bench/gossip_bw.py [--model gpt2]

Every rank owns a device arena, maps the others through CUDA IPC and runs the
all-to-all exchange of a model of the named size (BASELINE.json config 5 for
``gpt2``).  Times are CUDA events on the launching stream, max over ranks, after
warm-up; inputs exceed L2 for the large models, the small ones get an L2 flush.
Roofline = bytes that must enter each GPU over NVLink / 770 GB/s measured peer
bandwidth (B200_PROFILING.md); the fraction against the nominal 900 GB/s is printed too.
Prints one JSON line per (model, wire dtype) for the fused kernel, and one line per model for
the other data-plane variants: the copy-engine push + reduce (mode B through cp.async.bulk),
the remote-reduce mode A (cp.reduce.async.bulk), the two-shot in-switch all-reduce (NVLS,
multimem.ld_reduce - the upper bound for "every GPU trains"), and the model store to all peers
(init / diffusion): receiver pulls vs one bulk store per peer vs one multimem.st for all.
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
NVLINK_NOMINAL_GBS = 900.0


def timed(fn, prep, iters, warmup, dist, flush):
    """CUDA-event time of fn() (mean, best) over `iters`, max over ranks; prep() runs untimed before each."""
    ts = []
    for it in range(warmup + iters):
        prep()
        flush.zero_()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(it)
        e1.record()
        torch.cuda.synchronize()
        if it >= warmup:
            ts.append(e0.elapsed_time(e1))
    t = torch.tensor([sum(ts) / len(ts), min(ts)], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0]), float(t[1])


def variants(name, n, rank, local, world, args, dist, flush):
    """The data-plane variants besides the fused kernel, fp32 unless noted; one JSON line (rank 0)."""
    from p2pfl_b200 import ops
    from p2pfl_b200.parallel import multicast as mc
    from p2pfl_b200.parallel.arena import DeviceArena, PeerArena

    dev = torch.device(f"cuda:{local}")
    others = [r for r in range(world) if r != rank]
    out = {"bench": "gossip_variants", "model": name, "elements": n, "n_gpus": world}
    torch.manual_seed(rank)
    model = torch.randn(n, device=dev)
    w16 = torch.empty(n, dtype=torch.bfloat16, device=dev)
    weights = [1.0 / world] * world
    status = torch.zeros(4, dtype=torch.int32, device=dev)
    gathered = [torch.empty_like(model) for _ in range(world)]
    dist.all_gather(gathered, model)
    expect = sum(g / world for g in gathered)

    for wire_bf16 in (True, False):
        arena = DeviceArena(n, slots=world, tile=16384, wire_bf16=wire_bf16, device=local)
        handles = [None] * world
        dist.all_gather_object(handles, (arena.ipc_handle(), local))
        peers = {r: PeerArena.open_ipc(h, arena.geo, local, d) for r, (h, d) in enumerate(handles) if r != rank}
        tag = "bf16" if wire_bf16 else "fp32"
        wb = 2 if wire_bf16 else 4
        roof = (world - 1) * n * wb / (NVLINK_GBS * 1e9) * 1e3

        # ---- mode B with the copy engine: bulk push into the inbox slots, then the reduce kernel
        def bulk_exchange(it, seq0=1000):
            seq = seq0 + it
            bank = seq & 1
            push = ops.make_push_table([peers[r].slot_ptr(rank, bank) for r in others], [peers[r].flag_ptr(rank, bank) for r in others], wire_bf16)
            red = ops.make_reduce_table([arena.slot_ptr(r, bank) for r in others], [arena.flag_ptr(r, bank) for r in others],
                                        [weights[r] for r in others], rank, weights[rank], wire_bf16, True)
            ops.gossip_push_bulk(arena.params, seq, push)
            ops.gossip_reduce(arena.params, w16, 16384, seq, red, None, int(20e9), status)

        mean, best = timed(bulk_exchange, lambda: arena.params.copy_(model), args.iters, args.warmup, dist, flush)
        ref = expect if not wire_bf16 else sum(g.bfloat16().float() / world for g in gathered)
        out[f"bulk_push_reduce_{tag}_ms"] = mean
        out[f"bulk_push_reduce_{tag}_ms_best"] = best
        out[f"bulk_push_reduce_{tag}_frac_770"] = roof / mean
        out[f"bulk_push_reduce_{tag}_frac_900"] = roof * NVLINK_GBS / NVLINK_NOMINAL_GBS / mean
        out[f"bulk_push_reduce_{tag}_err"] = float((arena.params - ref).abs().max())

        if not wire_bf16:
            # ---- mode A: the peers' copy engines add into my accumulator (the outbox region), I only finalise
            counts = arena.flags_tensor(0, 0)
            acc_ptrs = {r: peers[r].region_ptr("outbox") for r in others}
            cnt_ptrs = {r: peers[r].flag_ptr(0, 0) for r in others}

            def prep_a():
                arena.params.copy_(model)
                arena.outbox.zero_()
                counts.zero_()

            def mode_a(it):
                table = ops.make_push_table([acc_ptrs[r] for r in others], [cnt_ptrs[r] for r in others], False)
                ops.gossip_push_reduce_bulk(arena.params, weights[rank], table)
                ops.gossip_finalize_acc(arena.params, w16, arena.outbox, counts, len(others), weights[rank], int(20e9), status)

            # pushes read params while finalize rewrites them: keep the sender's source separate
            src_copy = model.clone()

            def mode_a_safe(it):
                table = ops.make_push_table([acc_ptrs[r] for r in others], [cnt_ptrs[r] for r in others], False)
                ops.gossip_push_reduce_bulk(src_copy, weights[rank], table)
                ops.gossip_finalize_acc(arena.params, w16, arena.outbox, counts, len(others), weights[rank], int(20e9), status)

            mean, best = timed(mode_a_safe, prep_a, args.iters, args.warmup, dist, flush)
            out["remote_reduce_fp32_ms"] = mean
            out["remote_reduce_fp32_ms_best"] = best
            out["remote_reduce_fp32_frac_770"] = roof / mean
            out["remote_reduce_fp32_frac_900"] = roof * NVLINK_GBS / NVLINK_NOMINAL_GBS / mean
            out["remote_reduce_fp32_err"] = float((arena.params - expect).abs().max())

            # ---- model store to every peer (init / diffusion, fp32): pull by each receiver vs one bulk store per peer
            def store_bulk(it):
                table = ops.make_push_table([peers[r].region_ptr("outbox") for r in others], [peers[r].flag_ptr(1, 0) for r in others], False)
                if rank == 0:
                    ops.gossip_push_bulk(src_copy, 5000 + it, table)

            mean, best = timed(store_bulk, lambda: None, args.iters, args.warmup, dist, flush)
            out["store_bulk_unicast_ms"] = mean
            out["store_bulk_unicast_GBps_out"] = (world - 1) * n * 4 / (mean * 1e-3) / 1e9

            def store_pull(it):
                if rank != 0:
                    src = tensor_view(peers[0].region_ptr("params"), n, local)
                    ops.fedavg_into([src], [1.0], arena.outbox, None)

            from p2pfl_b200.parallel.arena import tensor_from_ptr

            def tensor_view(ptr, n_, d):
                return tensor_from_ptr(ptr, n_, torch.float32, d)

            mean, best = timed(store_pull, lambda: None, args.iters, args.warmup, dist, flush)
            out["store_pull_ms"] = mean
        for p in peers.values():
            p.close()
        dist.barrier()
        arena.close()

    # ---- NVLS: two-shot all-reduce and multicast store
    out["multicast_supported"] = bool(mc.supported(local))
    if out["multicast_supported"]:
        try:
            group = mc.MulticastGroup.distributed(dist, local, n)
            b = group.bindings[0]
            params = model.clone()
            mean, best = timed(lambda it: b.allreduce(params, w16, rank, world, weights[rank], int(20e9), status),
                               lambda: params.copy_(model), args.iters, args.warmup, dist, flush)
            out["nvls_allreduce_fp32_ms"] = mean
            out["nvls_allreduce_fp32_ms_best"] = best
            out["nvls_allreduce_fp32_err"] = float((params - expect).abs().max())
            out["nvls_allreduce_vs_unicast_roofline_770"] = (world - 1) * n * 4 / (NVLINK_GBS * 1e9) * 1e3 / mean
            mean, best = timed(lambda it: b.store(model, 7000 + it) if rank == 0 else None, lambda: None, args.iters, args.warmup, dist, flush)
            out["store_multicast_ms"] = mean
            out["store_multicast_GBps_delivered"] = (world - 1) * n * 4 / (mean * 1e-3) / 1e9
            dist.barrier()
            group.close()
        except Exception as e:  # creation can be refused (fabric state, container limits): say so, fall back
            out["multicast_error"] = f"{type(e).__name__}: {e}"[:200]
    if rank == 0:
        print(json.dumps(out), flush=True)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--models", default="mlp,resnet18,gpt2")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--tile", type=int, default=16384)
    ap.add_argument("--fused-only", action="store_true")
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
                    "fraction_of_nominal_900GBps": roof_ms * NVLINK_GBS / NVLINK_NOMINAL_GBS / float(t[0]),
                    "nccl_allgather_avg_ms_mean": float(bt[0]), "nccl_allgather_avg_ms_best": float(bt[1]),
                    "speedup_vs_nccl": float(bt[0]) / float(t[0]),
                }), flush=True)
            for p in peers.values():
                p.close()
            dist.barrier()
            arena.close()
        if not args.fused_only:
            variants(name, n, rank, local, world, args, dist, flush)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
