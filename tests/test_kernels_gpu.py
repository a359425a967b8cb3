"""Numerics of every sm_100a kernel against a plain PyTorch fp32 reference."""

import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from p2pfl_b200 import ops as _ops

    _ops.lib()  # fail loudly when the library is missing on a GPU box
    return _ops


DEV = "cuda:0"


def test_library_loaded(ops):
    assert ops.lib().p2pfl_version() == 100
    assert torch.cuda.get_device_capability(0)[0] == 10


# ------------------------------------------------------------------ optimizers
@pytest.mark.parametrize("n", [32, 235168, 1000003 * 4])
def test_adam_matches_torch(ops, n):
    torch.manual_seed(0)
    p_ref = torch.randn(n, device=DEV)
    p = p_ref.clone()
    m = torch.zeros_like(p)
    v = torch.zeros_like(p)
    p16 = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    step = torch.zeros(1, device=DEV, dtype=torch.int32)
    param = torch.nn.Parameter(p_ref.clone())
    opt = torch.optim.Adam([param], lr=1e-3)
    for it in range(3):
        g = torch.randn(n, device=DEV)
        param.grad = g.clone()
        opt.step()
        ops.adam_step(p, g, m, v, p16, 1e-3, step, it + 1)
    torch.cuda.synchronize()
    assert torch.allclose(p, param.data, atol=1e-6, rtol=1e-5)
    assert torch.equal(p16, p.to(torch.bfloat16))


def test_adam_device_step_counter(ops):
    n = 4096
    p = torch.ones(n, device=DEV)
    q = p.clone()
    g = torch.full((n,), 0.5, device=DEV)
    m1, v1, m2, v2 = (torch.zeros(n, device=DEV) for _ in range(4))
    step = torch.zeros(1, device=DEV, dtype=torch.int32)
    for it in range(4):
        ops.adam_step(p, g, m1, v1, None, 1e-2, step, 1)  # t = *step + 1
        ops.add_i32(step, 1)
        ops.adam_step(q, g, m2, v2, None, 1e-2, None, it + 1)
    torch.cuda.synchronize()
    assert int(step.item()) == 4
    assert torch.equal(p, q)


def test_sgd(ops):
    n = 10007 * 8
    p = torch.randn(n, device=DEV)
    ref = p.clone()
    g = torch.randn(n, device=DEV)
    mom = torch.zeros(n, device=DEV)
    p16 = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    ops.sgd_step(p, g, mom, p16, 0.1, 0.9)
    ops.sgd_step(p, g, mom, p16, 0.1, 0.9)
    buf = g.clone()
    ref -= 0.1 * buf
    buf = 0.9 * buf + g
    ref -= 0.1 * buf
    assert torch.allclose(p, ref, atol=1e-6)
    assert torch.equal(p16, p.to(torch.bfloat16))


def test_cast_and_transpose(ops):
    src = torch.randn(300 * 70 + 64 + 10 * 128, device=DEV)
    dst = torch.zeros(4096 * 16, device=DEV, dtype=torch.bfloat16)
    # two matrices: 300x70 at 0 -> ld 304 ; 10x128 at 21064 -> ld 16
    descs = [(0, 0, 300, 70, 304), (300 * 70 + 64, 70 * 304, 10, 128, 16)]
    ops.transpose_cast(src, dst, descs)
    a = src[: 300 * 70].view(300, 70)
    b = src[300 * 70 + 64 :].view(10, 128)
    at = dst[: 70 * 304].view(70, 304)
    bt = dst[70 * 304 : 70 * 304 + 128 * 16].view(128, 16)
    assert torch.equal(at[:, :300], a.t().to(torch.bfloat16))
    assert torch.equal(bt[:, :10], b.t().to(torch.bfloat16))
    assert float(at[:, 300:].abs().sum()) == 0
    c = torch.empty(src.numel(), device=DEV, dtype=torch.bfloat16)
    ops.cast_f32_bf16(src, c)
    assert torch.equal(c, src.to(torch.bfloat16))


def test_gather_rows(ops):
    n, feat, batch = 100, 784, 32
    src = torch.randn(n, feat, device=DEV)
    labels = torch.randint(0, 10, (n,), device=DEV, dtype=torch.int32)
    perm = torch.randperm(n, device=DEV).to(torch.int32)
    nb = (n + batch - 1) // batch
    dst = torch.zeros(n, feat, device=DEV, dtype=torch.bfloat16)
    dst_t = torch.zeros(nb, feat, batch, device=DEV, dtype=torch.bfloat16)
    lab = torch.zeros(n, device=DEV, dtype=torch.int32)
    ops.gather_rows(src, perm, n, feat, batch, dst, dst_t, labels, lab)
    ref = src[perm.long()].to(torch.bfloat16)
    assert torch.equal(dst, ref)
    assert torch.equal(lab, labels[perm.long()])
    for b in range(nb):
        rows = ref[b * batch : (b + 1) * batch]
        assert torch.equal(dst_t[b][:, : rows.shape[0]], rows.t())


# ---------------------------------------------------------------------- losses
@pytest.mark.parametrize("B,Cn,ld", [(32, 10, 16), (32, 10, 10), (100, 1000, 1000), (7, 50, 64)])
def test_softmax_xent_fwd_bwd(ops, B, Cn, ld):
    torch.manual_seed(1)
    logits = torch.zeros(B, ld, device=DEV)
    logits[:, :Cn] = torch.randn(B, Cn, device=DEV) * 3
    labels = torch.randint(0, Cn, (B,), device=DEV, dtype=torch.int32)
    ldd = (Cn + 7) // 8 * 8
    ldt = (B + 7) // 8 * 8
    dl = torch.zeros(B, ldd, device=DEV, dtype=torch.bfloat16)
    dlt = torch.zeros(Cn, ldt, device=DEV, dtype=torch.bfloat16)
    db = torch.zeros(Cn, device=DEV)
    loss = torch.zeros(1, device=DEV)
    ops.softmax_xent(logits, ld, labels, B, Cn, dl, ldd, dlt, ldt, db, loss)
    x = logits[:, :Cn].clone().requires_grad_(True)
    ref = F.cross_entropy(x, labels.long(), reduction="mean")
    ref.backward()
    assert math.isclose(float(loss) / B, float(ref), rel_tol=1e-5, abs_tol=1e-6)
    assert torch.allclose(dl[:, :Cn].float(), x.grad, atol=2e-3 / B + 1e-4, rtol=1e-2)
    assert torch.equal(dlt[:, :B], dl[:, :Cn].t())
    assert torch.allclose(db, dl[:, :Cn].float().sum(0), atol=1e-4)
    assert float(dl[:, Cn:].abs().sum()) == 0


def test_eval_reduce(ops):
    B, Cn = 1250, 10
    logits = torch.randn(B, 16, device=DEV)
    labels = torch.randint(0, Cn, (B,), device=DEV, dtype=torch.int32)
    loss = torch.zeros(1, device=DEV)
    correct = torch.zeros(1, device=DEV, dtype=torch.int32)
    ops.eval_reduce(logits, 16, labels, B, Cn, loss, correct)
    ref_loss = F.cross_entropy(logits[:, :Cn], labels.long(), reduction="sum")
    ref_correct = (logits[:, :Cn].argmax(1) == labels.long()).sum()
    assert math.isclose(float(loss), float(ref_loss), rel_tol=1e-5)
    assert int(correct) == int(ref_correct)


# ---------------------------------------------------------------------- FedAvg
@pytest.mark.parametrize("k", [1, 3, 8, 20])
def test_fedavg_kernel(ops, k):
    n = 235168
    torch.manual_seed(k)
    models = [torch.randn(n, device=DEV) for _ in range(k)]
    if k >= 3:
        models[1] = models[1].to(torch.bfloat16)  # mixed dtypes
    w = torch.rand(k).tolist()
    tot = sum(w)
    w = [x / tot for x in w]
    out = ops.fedavg_flat(models, w)
    ref = torch.zeros(n, device=DEV)
    for m, x in zip(models, w):
        ref += m.float() * x
    assert torch.allclose(out, ref, atol=1e-5)
    out16 = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    ops.fedavg_into(models, w, None, out16)
    assert torch.allclose(out16.float(), ref, atol=2e-2, rtol=1e-2)


def test_fedavg_flat_path_matches_cpu_oracle(ops):
    from p2pfl_b200.learning.aggregators.fedavg import FedAvg, fedavg_state_dicts
    from p2pfl_b200.learning.torch_learner import TorchLearner
    from p2pfl_b200.models import MLP

    learners = [TorchLearner(MLP(seed=s), None, device=DEV) for s in range(3)]
    models = {f"n{i}": (l.export_model(), 100 + 50 * i) for i, l in enumerate(learners)}
    out = FedAvg().aggregate(models)
    oracle = fedavg_state_dicts({k: ({n: t.cpu() for n, t in m.state_dict().items()}, w) for k, (m, w) in models.items()})
    for name, t in out.state_dict().items():
        assert torch.allclose(t.cpu(), oracle[name], atol=1e-6)


# ------------------------------------------------------------------------ GEMM
def _gemm_ref(A, B, bias, bias_per_row, relu, mask, alpha):
    D = alpha * (A.float() @ B.float().t())
    if bias is not None:
        D = D + (bias[:, None] if bias_per_row else bias[None, :])
    if relu:
        D = torch.relu(D)
    if mask is not None:
        D = torch.where(mask.float() > 0, D, torch.zeros_like(D))
    return D


GEMM_SHAPES = [
    # (M, N, K)   MLP forward / dgrad / wgrad shapes (swap-AB) and bigger ones
    (256, 32, 784),
    (128, 32, 256),
    (10, 32, 128),
    (128, 32, 10),
    (256, 32, 128),
    (256, 784, 32),
    (10, 128, 32),
    (2048, 32, 3136),
    (128, 128, 64),
    (300, 200, 136),
    (512, 512, 512),
    (1024, 256, 4096),
    # many more tiles than SMs: persistent kernel with double-buffered TMEM accumulators
    (4096, 4096, 256),
    (5000, 3000, 136),
]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_tn_plain(ops, M, N, K):
    torch.manual_seed(M * 7 + N * 3 + K)
    lda = (K + 7) // 8 * 8
    A = torch.zeros(M, lda, device=DEV, dtype=torch.bfloat16)
    B = torch.zeros(N, lda, device=DEV, dtype=torch.bfloat16)
    A[:, :K] = (torch.randn(M, K, device=DEV) / math.sqrt(K)).to(torch.bfloat16)
    B[:, :K] = torch.randn(N, K, device=DEV).to(torch.bfloat16)
    out = torch.full((M, N), float("nan"), device=DEV)
    ops.gemm_tn(A, lda, B, lda, M, N, K, out_f32=out, ld_f32=N)
    torch.cuda.synchronize()
    ref = _gemm_ref(A[:, :K], B[:, :K], None, True, False, None, 1.0)
    err = (out - ref).abs().max().item()
    assert torch.allclose(out, ref, atol=2e-3, rtol=2e-3), f"max abs err {err}"


@pytest.mark.parametrize("M,N,K", [(256, 32, 784), (128, 96, 200), (10, 32, 128)])
def test_gemm_tn_epilogue_all_outputs(ops, M, N, K):
    torch.manual_seed(5)
    lda = (K + 7) // 8 * 8
    A = torch.zeros(M, lda, device=DEV, dtype=torch.bfloat16)
    B = torch.zeros(N, lda, device=DEV, dtype=torch.bfloat16)
    A[:, :K] = (torch.randn(M, K, device=DEV) / math.sqrt(K)).to(torch.bfloat16)
    B[:, :K] = torch.randn(N, K, device=DEV).to(torch.bfloat16)
    bias = torch.randn(M, device=DEV)
    ldn = (N + 7) // 8 * 8
    ldm = (M + 7) // 8 * 8
    mask = torch.randn(M, ldn, device=DEV).to(torch.bfloat16)
    o32 = torch.zeros(M, ldn, device=DEV)
    o16 = torch.zeros(M, ldn, device=DEV, dtype=torch.bfloat16)
    o32t = torch.zeros(N, ldm, device=DEV)
    o16t = torch.zeros(N, ldm, device=DEV, dtype=torch.bfloat16)
    rs = torch.zeros(M, device=DEV)
    ops.gemm_tn(
        A, lda, B, lda, M, N, K, bias=bias, bias_per_row=True, relu=True, mask=mask, ld_mask=ldn,
        out_f32=o32, ld_f32=ldn, out_bf16=o16, ld_bf16=ldn, out_f32_t=o32t, ld_f32_t=ldm,
        out_bf16_t=o16t, ld_bf16_t=ldm, rowsum=rs, alpha=0.5,
    )  # fmt: skip
    ref = _gemm_ref(A[:, :K], B[:, :K], bias, True, True, mask[:, :N], 0.5)
    assert torch.allclose(o32[:, :N], ref, atol=2e-3, rtol=2e-3)
    assert torch.allclose(o16[:, :N].float(), ref, atol=2e-2, rtol=2e-2)
    assert torch.allclose(o32t[:, :M], ref.t(), atol=2e-3, rtol=2e-3)
    assert torch.allclose(o16t[:, :M].float(), ref.t(), atol=2e-2, rtol=2e-2)
    assert torch.allclose(rs, ref.sum(1), atol=1e-2, rtol=1e-2)
    assert float(o32[:, N:].abs().sum()) == 0 and float(o32t[:, M:].abs().sum()) == 0


def test_gemm_tn_bias_per_column(ops):
    M, N, K = 64, 48, 72
    A = (torch.randn(M, K, device=DEV) / 8).to(torch.bfloat16)
    B = torch.randn(N, K, device=DEV).to(torch.bfloat16)
    bias = torch.randn(N, device=DEV)
    out = torch.zeros(M, N, device=DEV)
    ops.gemm_tn(A, K, B, K, M, N, K, bias=bias, bias_per_row=False, out_f32=out, ld_f32=N)
    ref = _gemm_ref(A, B, bias, False, False, None, 1.0)
    assert torch.allclose(out, ref, atol=2e-3, rtol=2e-3)


# ---------------------------------------------------------------------- gossip
@pytest.mark.parametrize("wire_bf16", [True, False])
@pytest.mark.parametrize("ranks", [2, 4, 8])
def test_gossip_fused_emulated(ops, ranks, wire_bf16):
    """R logical ranks exchange and FedAvg in ONE cooperative launch; the result must
    equal the CPU FedAvg oracle and be bit-identical on every rank."""
    from p2pfl_b200.parallel.arena import emulate_exchange

    n, tile = 235168, 16384
    torch.manual_seed(ranks)
    models = [torch.randn(n, device=DEV) for _ in range(ranks)]
    samples = [100 + 37 * r for r in range(ranks)]
    outs, status = emulate_exchange(models, samples, tile=tile, wire_bf16=wire_bf16, seq=1)
    tot = float(sum(samples))
    ref = torch.zeros(n, device=DEV)
    for m, s in zip(models, samples):
        mm = m.to(torch.bfloat16).float() if wire_bf16 else m
        ref += mm * (s / tot)
    assert status[:, 0].sum().item() == 0, "a rank timed out"
    for r in range(ranks):
        assert torch.allclose(outs[r], ref, atol=1e-5), f"rank {r}: {(outs[r]-ref).abs().max().item()}"
        assert torch.equal(outs[r], outs[0]), "ranks disagree bit-wise"


def test_gossip_push_then_reduce_and_timeout(ops):
    """Split mode (G1 then G2) on one device, plus the abort path: a contributor that
    never arrives is dropped and the tile is renormalised."""
    from p2pfl_b200.parallel.arena import DeviceArena

    n, tile = 65536 + 2048, 16384
    a = DeviceArena(n, slots=2, tile=tile, wire_bf16=True, device=0)
    b = DeviceArena(n, slots=2, tile=tile, wire_bf16=True, device=0)
    try:
        torch.manual_seed(0)
        a.params.copy_(torch.randn(n, device=DEV))
        b.params.copy_(torch.randn(n, device=DEV))
        pa, pb = a.params.clone(), b.params.clone()
        # a pushes into b's slot 0
        push = ops.make_push_table([b.slot_ptr(0, bank=1)], [b.flag_ptr(0, bank=1)], True)
        ops.gossip_push(a.params, tile, 1, push)
        red = ops.make_reduce_table([b.slot_ptr(0, bank=1)], [b.flag_ptr(0, bank=1)], [0.25], 1, 0.75, True, True)
        status = torch.zeros(4, device=DEV, dtype=torch.int32)
        ops.gossip_reduce(b.params, None, tile, 1, red, None, 0, status)
        torch.cuda.synchronize()
        ref = 0.25 * pa.to(torch.bfloat16).float() + 0.75 * pb.to(torch.bfloat16).float()
        assert torch.allclose(b.params, ref, atol=1e-6)
        assert status[0].item() == 0 and status[1].item() == b.num_tiles
        # now expect a second contributor (slot 1) that never pushes: bounded wait, then renormalise
        b.params.copy_(pb)
        ops.gossip_push(a.params, tile, 3, ops.make_push_table([b.slot_ptr(0, bank=1)], [b.flag_ptr(0, bank=1)], True))
        red2 = ops.make_reduce_table(
            [b.slot_ptr(0, bank=1), b.slot_ptr(1, bank=1)], [b.flag_ptr(0, bank=1), b.flag_ptr(1, bank=1)],
            [0.25, 0.25], 2, 0.5, True, True,
        )  # fmt: skip
        status.zero_()
        ops.gossip_reduce(b.params, None, tile, 3, red2, None, 2_000_000, status)
        torch.cuda.synchronize()
        ref2 = (0.25 * pa.to(torch.bfloat16).float() + 0.5 * pb.to(torch.bfloat16).float()) / 0.75
        assert status[0].item() == 1
        assert torch.allclose(b.params, ref2, atol=1e-5)
    finally:
        a.close()
        b.close()


def test_native_watchdog(ops):
    """``p2pfl_rt_watch`` (csrc/watchdog.cu): returns with the event; on a timeout or a cancel word set
    by another thread it flips the device abort word itself and the waiting kernel leaves (renormalised
    result, status says incomplete); a slice hands back while the kernel is still waiting."""
    import ctypes
    import threading
    import time

    from p2pfl_b200.parallel.arena import DeviceArena

    n, tile = 65536, 16384
    b = DeviceArena(n, slots=2, tile=tile, wire_bf16=True, device=0)
    try:
        ev = torch.cuda.Event()
        ev.record()
        verdict, _ = ops.rt_watch(0, ev, None, None, 1.0, 0.0)
        assert verdict == ops.WATCH_DONE
        budget = int(30 * ops.rt_sm_clock_hz(0))  # the kernel would give up by itself after 30 s
        for k, mode in enumerate(("timeout", "cancel", "slice")):
            torch.manual_seed(k)
            own = torch.randn(n, device=DEV)
            b.params.copy_(own)
            b.abort.zero_()
            b.status.zero_()
            seq = 5 + 2 * k  # nobody ever pushes for it
            red = ops.make_reduce_table([b.slot_ptr(0, bank=1)], [b.flag_ptr(0, bank=1)], [0.5], 1, 0.5, True, True)
            ops.gossip_reduce(b.params, None, tile, seq, red, b.abort, budget, b.status)
            ev = torch.cuda.Event()
            ev.record()
            cancel = ctypes.c_int32(0)
            t0 = time.time()
            if mode == "timeout":
                verdict, waited = ops.rt_watch(0, ev, b.abort, cancel, 0.1, 0.0, 5.0)
                assert verdict == ops.WATCH_TIMEOUT and 0.09 < waited < 3.0
            elif mode == "cancel":
                threading.Timer(0.05, lambda: setattr(cancel, "value", 1)).start()
                verdict, waited = ops.rt_watch(0, ev, b.abort, cancel, 20.0, 0.0, 5.0)
                assert verdict == ops.WATCH_CANCELLED and 0.04 < waited < 3.0
            else:
                verdict, waited = ops.rt_watch(0, ev, b.abort, cancel, 20.0, 0.01, 5.0)
                assert verdict == ops.WATCH_SLICE and 0.009 < waited < 1.0
                assert not ev.query()  # still waiting for its contributor
                cancel.value = 1
                verdict, _ = ops.rt_watch(0, ev, b.abort, cancel, 20.0, 0.01, 5.0)
                assert verdict == ops.WATCH_CANCELLED
            assert time.time() - t0 < 6.0
            assert ev.query()
            torch.cuda.synchronize()
            assert int(b.abort.item()) == 1 and int(b.status[0].item()) == 1
            # the missing contributor was dropped and the tile renormalised: the own model (bf16-rounded)
            assert torch.allclose(b.params, own.to(torch.bfloat16).float(), atol=1e-5)
    finally:
        b.close()


# ------------------------------------------------- launch-reduction variants
def test_gemm_grouped_matches_single_launches(ops):
    """A layer's wgrad and dgrad in ONE launch (different shapes, tile widths, epilogues)."""
    torch.manual_seed(11)
    B = 32
    dZT = (torch.randn(128, B, device=DEV) / 8).to(torch.bfloat16)  # [out, B]
    HT = torch.randn(256, B, device=DEV).to(torch.bfloat16)  # [in, B]
    wT = (torch.randn(256, 128, device=DEV) / 16).to(torch.bfloat16)  # [in, out]
    dZ = dZT.t().contiguous()  # [B, out]
    grad_w = torch.full((128, 256), 7.0, device=DEV)
    dHT = torch.zeros(256, B, device=DEV, dtype=torch.bfloat16)
    dH = torch.zeros(B, 256, device=DEV, dtype=torch.bfloat16)
    db = torch.full((256,), 5.0, device=DEV)  # stale content must be overwritten (N <= 32 stores)
    ops.gemm_tn_grouped(
        torch.device(DEV),
        [
            ops.gemm_args(dZT, B, HT, B, 128, 256, B, out_f32=grad_w, ld_f32=256),
            ops.gemm_args(wT, 128, dZ, 128, 256, B, 128, mask=HT, ld_mask=B, out_bf16=dHT, ld_bf16=B,
                          out_bf16_t=dH, ld_bf16_t=256, rowsum=db),
        ],
    )  # fmt: skip
    ref_w = dZT.float() @ HT.float().t()
    ref_d = torch.where(HT.float() > 0, wT.float() @ dZ.float().t(), torch.zeros(256, B, device=DEV))
    assert torch.allclose(grad_w, ref_w, atol=2e-3, rtol=2e-3)
    assert torch.allclose(dHT.float(), ref_d, atol=2e-2, rtol=2e-2)
    assert torch.equal(dH, dHT.t())
    assert torch.allclose(db, ref_d.sum(1), atol=1e-2, rtol=1e-2)


def test_adam_with_transposed_copies(ops):
    n = 300 * 70 + 64 + 10 * 128
    torch.manual_seed(2)
    p = torch.randn(n, device=DEV)
    ref = p.clone()
    g = torch.randn(n, device=DEV)
    m, v, m2, v2 = (torch.zeros(n, device=DEV) for _ in range(4))
    p16 = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    wt = torch.zeros(70 * 304 + 128 * 16, device=DEV, dtype=torch.bfloat16)
    descs = [(0, 0, 300, 70, 304), (300 * 70 + 64, 70 * 304, 10, 128, 16)]
    ops.adam_step_t(p, g, m, v, p16, 1e-2, None, 1, wt, descs)
    ops.adam_step(ref, g, m2, v2, None, 1e-2, None, 1)
    assert torch.equal(p, ref) and torch.equal(p16, p.to(torch.bfloat16))
    a = p[: 300 * 70].view(300, 70)
    b = p[300 * 70 + 64 :].view(10, 128)
    assert torch.equal(wt[: 70 * 304].view(70, 304)[:, :300], a.t().to(torch.bfloat16))
    assert torch.equal(wt[70 * 304 :].view(128, 16)[:, :10], b.t().to(torch.bfloat16))


def test_softmax_xent_overwrites_dbias(ops):
    B, Cn = 32, 10
    logits = torch.randn(B, 16, device=DEV)
    labels = torch.randint(0, Cn, (B,), device=DEV, dtype=torch.int32)
    dl = torch.zeros(B, 16, device=DEV, dtype=torch.bfloat16)
    for big in (False, True):
        Bb = 512 if big else B  # > 256 rows takes the multi-block path
        lg = torch.randn(Bb, 16, device=DEV)
        lb = torch.randint(0, Cn, (Bb,), device=DEV, dtype=torch.int32)
        d = torch.zeros(Bb, 16, device=DEV, dtype=torch.bfloat16)
        db = torch.full((Cn,), 123.0, device=DEV)
        ops.softmax_xent(lg, 16, lb, Bb, Cn, d, 16, None, 0, db, None)
        assert torch.allclose(db, d[:, :Cn].float().sum(0), atol=1e-4), big


# ----------------------------------------------- MN-major operands and split-K
@pytest.mark.parametrize("a_mn,b_mn", [(True, True), (False, True), (True, False)])
@pytest.mark.parametrize("M,N,K", [(64, 576, 4096), (128, 27, 1000), (512, 4608, 1024), (200, 136, 64), (64, 64, 65536)])
def test_gemm_mn_major_operands(ops, M, N, K, a_mn, b_mn):
    """Operands stored [K, MN] (MN contiguous) are read in place: no transposed copies."""
    torch.manual_seed(M + N + K)
    A = (torch.randn(M, K, device=DEV) / math.sqrt(K)).to(torch.bfloat16)
    B = torch.randn(N, K, device=DEV).to(torch.bfloat16)
    ref = A.float() @ B.float().t()
    ldm, ldn, ldk = (M + 7) // 8 * 8, (N + 7) // 8 * 8, (K + 7) // 8 * 8
    if a_mn:
        a_op = torch.zeros(K, ldm, device=DEV, dtype=torch.bfloat16)
        a_op[:, :M] = A.t()
        lda = ldm
    else:
        a_op = torch.zeros(M, ldk, device=DEV, dtype=torch.bfloat16)
        a_op[:, :K] = A
        lda = ldk
    if b_mn:
        b_op = torch.zeros(K, ldn, device=DEV, dtype=torch.bfloat16)
        b_op[:, :N] = B.t()
        ldb = ldn
    else:
        b_op = torch.zeros(N, ldk, device=DEV, dtype=torch.bfloat16)
        b_op[:, :K] = B
        ldb = ldk
    out = torch.full((M, N), float("nan"), device=DEV)
    ops.gemm_tn(a_op, lda, b_op, ldb, M, N, K, out_f32=out, ld_f32=N, a_mn=a_mn, b_mn=b_mn)
    torch.cuda.synchronize()
    err = (out - ref).abs().max().item()
    assert torch.allclose(out, ref, atol=3e-3, rtol=3e-3), f"max abs err {err}"


@pytest.mark.parametrize("split", [-1, 3, 16])
@pytest.mark.parametrize("M,N,K,mn", [(64, 576, 65536, True), (64, 27, 8192, True), (128, 256, 4096, False), (512, 4608, 1024, True)])
def test_gemm_split_k_accumulates(ops, M, N, K, mn, split):
    """Weight-gradient shape: small output, long reduction, K slices add up atomically."""
    torch.manual_seed(split + M)
    A = (torch.randn(M, K, device=DEV) / math.sqrt(K)).to(torch.bfloat16)
    B = torch.randn(N, K, device=DEV).to(torch.bfloat16)
    ref = A.float() @ B.float().t()
    ld = (N + 7) // 8 * 8
    out = torch.zeros(M, ld, device=DEV)
    if mn:
        a_op, b_op = A.t().contiguous(), torch.zeros(K, ld, device=DEV, dtype=torch.bfloat16)
        b_op[:, :N] = B.t()
        ops.gemm_tn(a_op, M, b_op, ld, M, N, K, out_f32=out, ld_f32=ld, split_k=split, a_mn=True, b_mn=True)
    else:
        ops.gemm_tn(A, K, B, K, M, N, K, out_f32=out, ld_f32=ld, split_k=split)
    torch.cuda.synchronize()
    err = (out[:, :N] - ref).abs().max().item()
    assert torch.allclose(out[:, :N], ref, atol=3e-3, rtol=3e-3), f"max abs err {err}"
    assert float(out[:, N:].abs().max()) == 0.0 if ld > N else True


def test_gemm_split_k_rejects_fused_epilogue(ops):
    A = torch.zeros(64, 4096, device=DEV, dtype=torch.bfloat16)
    out = torch.zeros(64, 64, device=DEV, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="split-K"):
        ops.gemm_tn(A, 4096, A, 4096, 64, 64, 4096, out_bf16=out, ld_bf16=64, split_k=4)


def test_gemm_grouped_conv_backward_pair(ops):
    """The ResNet backward launch: split-K MN-major wgrad + MN-major-B dgrad in one grid."""
    torch.manual_seed(3)
    pix, cout, Kc = 4096, 128, 1152
    dz = (torch.randn(pix, cout, device=DEV) / 16).to(torch.bfloat16)
    col = torch.randn(pix, Kc, device=DEV).to(torch.bfloat16)
    w = (torch.randn(cout, Kc, device=DEV) / 32).to(torch.bfloat16)
    gw = torch.zeros(cout, Kc, device=DEV)
    dcol = torch.zeros(pix, Kc, device=DEV, dtype=torch.bfloat16)
    ops.gemm_tn_grouped(
        torch.device(DEV),
        [
            ops.gemm_args(dz, cout, col, Kc, cout, Kc, pix, out_f32=gw, ld_f32=Kc, split_k=-1, a_mn=True, b_mn=True),
            ops.gemm_args(dz, cout, w, Kc, pix, Kc, cout, out_bf16=dcol, ld_bf16=Kc, b_mn=True),
        ],
    )
    torch.cuda.synchronize()
    ref_w = dz.float().t() @ col.float()
    ref_d = dz.float() @ w.float()
    assert torch.allclose(gw, ref_w, atol=2e-2, rtol=1e-2), (gw - ref_w).abs().max().item()
    assert torch.allclose(dcol.float(), ref_d, atol=2e-2, rtol=2e-2), (dcol.float() - ref_d).abs().max().item()


@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (True, True)])
def test_gemm_persistent_with_mn_major_operands(ops, a_mn, b_mn):
    """Many more tiles than SMs (the layer-1 data gradient: M = pixels, K = Cout = 64)."""
    torch.manual_seed(9)
    M, N, K = 65536, 576, 64
    A = (torch.randn(M, K, device=DEV) / 8).to(torch.bfloat16)
    W = torch.randn(K, N, device=DEV).to(torch.bfloat16)  # [K, N]: MN-major B
    a_op, lda = (A.t().contiguous(), M) if a_mn else (A, K)
    out = torch.zeros(M, N, device=DEV, dtype=torch.bfloat16)
    ops.gemm_tn(a_op, lda, W, N, M, N, K, out_bf16=out, ld_bf16=N, a_mn=a_mn, b_mn=b_mn)
    torch.cuda.synchronize()
    ref = A.float() @ W.float()
    assert torch.allclose(out.float(), ref, atol=3e-2, rtol=2e-2), (out.float() - ref).abs().max().item()
