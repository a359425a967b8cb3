"""The library-collective baseline of the round's exchange (communication/nvlink/collective_exchange.py),
on CPU tensors over gloo: two processes, one node each."""

import os
import socket
import sys
from types import SimpleNamespace

import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank(rank: int, world: int, port: int, out) -> None:
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from p2pfl_b200.communication.memory import InMemoryCommunicationProtocol
    from p2pfl_b200.communication.nvlink.collective_exchange import with_collective_exchange

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(rank)
        flat = torch.randn(1000)
        mine = flat.clone()
        n = 100 * (rank + 1)
        written = []
        learner = SimpleNamespace(flat=flat, get_num_samples=lambda: (n, 10), _params_written=lambda: written.append(1))
        proto = with_collective_exchange(InMemoryCommunicationProtocol)(f"node-{rank}")
        addrs = [f"node-{r}" for r in range(world)]
        state = SimpleNamespace(addr=f"node-{rank}", train_set=addrs, round=0)
        plan = proto.plan_fused_exchange(state, learner)
        assert plan is not None and plan.can_prelaunch
        res = plan.run(7)
        gathered = [torch.empty(1000) for _ in range(world)]
        dist.all_gather(gathered, mine)
        want = sum(100 * (r + 1) * g for r, g in enumerate(gathered)) / sum(100 * (r + 1) for r in range(world))
        ok = torch.allclose(flat, want, atol=1e-6) and res == (addrs, sum(100 * (r + 1) for r in range(world))) and written == [1]
        # the other ranks must hold the same bits
        both = [torch.empty(1000) for _ in range(world)]
        dist.all_gather(both, flat)
        ok = ok and all(torch.equal(both[0], b) for b in both)
        # a train set that is not "every rank" gets no plan
        state2 = SimpleNamespace(addr=f"node-{rank}", train_set=addrs[:1], round=0)
        ok = ok and proto.plan_fused_exchange(state2, learner) is None
        out.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_all_gather_average_matches_fedavg_and_is_identical_on_every_rank():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(out.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(30)
    assert res == [(0, True), (1, True)], res
