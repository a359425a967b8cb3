"""The C++ oracles against PyTorch on the CPU (so that a kernel that disagrees with an oracle is
the kernel's fault), and - on a GPU - the kernels against the oracles."""

import pytest
import torch
import torch.nn.functional as F


@pytest.fixture(scope="module")
def oracle():
    import build

    build.build_oracle()
    from p2pfl_b200.ops import oracle as o

    return o


def test_oracles_agree_with_pytorch(oracle):
    torch.manual_seed(0)
    A, B, bias = torch.randn(37, 100), torch.randn(19, 100), torch.randn(19)
    assert torch.allclose(oracle.gemm_tn(A, B, bias, relu=True), F.relu(A @ B.t() + bias), atol=1e-4)
    logits, y = torch.randn(32, 10) * 3, torch.randint(0, 10, (32,))
    loss, d = oracle.softmax_xent(logits, y)
    lg = logits.clone().requires_grad_(True)
    ref = F.cross_entropy(lg, y)
    ref.backward()
    assert abs(loss / 32 - float(ref)) < 1e-5 and torch.allclose(d, lg.grad, atol=1e-6)
    el, correct = oracle.evaluate(logits, y)
    assert abs(el - loss) < 1e-9 and correct == int((logits.argmax(1) == y).sum())
    p = torch.randn(1000, requires_grad=True)
    opt = torch.optim.Adam([p], lr=1e-3)
    pm, m, v = p.detach().clone(), torch.zeros(1000), torch.zeros(1000)
    for t in range(1, 4):
        g = torch.randn(1000)
        p.grad = g.clone()
        opt.step()
        pm, m, v = oracle.adam(pm, g, m, v, 1e-3, t)
    assert torch.allclose(pm, p.detach(), atol=1e-6)
    models, w = [torch.randn(5000) for _ in range(4)], [0.1, 0.2, 0.3, 0.4]
    assert torch.allclose(oracle.fedavg(models, w), sum(wi * mi for wi, mi in zip(w, models)), atol=1e-6)
    x, wt, b = torch.randn(2, 3, 12, 12), torch.randn(5, 3, 3, 3), torch.randn(5)
    assert torch.allclose(oracle.conv2d(x, wt, b, stride=2, pad=1), F.conv2d(x, wt, b, stride=2, padding=1), atol=1e-4)
    assert torch.equal(oracle.maxpool2(x), F.max_pool2d(x, 2))
    gamma, beta = torch.rand(3) + 0.5, torch.randn(3)
    ref_bn = F.relu(F.batch_norm(x, None, None, gamma, beta, training=True, eps=1e-5))
    assert torch.allclose(oracle.batchnorm(x, gamma, beta, relu=True), ref_bn, atol=1e-5)


@pytest.mark.gpu
def test_kernels_agree_with_the_cpp_oracles(oracle):
    from p2pfl_b200 import ops

    dev = "cuda:0"
    torch.manual_seed(1)
    # GEMM + bias + ReLU (bf16 operands, fp32 accumulate)
    A = (torch.randn(96, 256) / 8).bfloat16()
    B = (torch.randn(80, 256) / 8).bfloat16()
    bias = torch.randn(80)
    out = torch.full((96, 80), float("nan"), device=dev)
    ops.gemm_tn(A.to(dev), 256, B.to(dev), 256, 96, 80, 256, bias=bias.to(dev), bias_per_row=False, relu=True, out_f32=out, ld_f32=80)
    torch.cuda.synchronize()
    assert torch.allclose(out.cpu(), oracle.gemm_tn(A.float(), B.float(), bias, relu=True), atol=2e-3, rtol=1e-3)
    # FedAvg
    models = [torch.randn(70000, device=dev) for _ in range(3)]
    w = [0.5, 0.3, 0.2]
    got = torch.empty(70000, device=dev)
    ops.fedavg_into(models, w, got, None)
    torch.cuda.synchronize()
    assert torch.allclose(got.cpu(), oracle.fedavg(models, w), atol=1e-6)
    # Adam (two steps)
    p = torch.randn(50000, device=dev)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    step = torch.zeros(1, dtype=torch.int32, device=dev)
    po, mo, vo = p.cpu(), m.cpu(), v.cpu()
    for t in (1, 2):
        g = torch.randn(50000, device=dev)
        ops.adam_step(p, g, m, v, None, 1e-3, step, t)
        po, mo, vo = oracle.adam(po, g, mo, vo, 1e-3, t)
    torch.cuda.synchronize()
    assert torch.allclose(p.cpu(), po, atol=1e-6) and torch.allclose(v.cpu(), vo, atol=1e-7)
    # softmax cross-entropy forward + backward, and the evaluation reduce
    logits = (torch.randn(32, 16, device=dev) * 2).contiguous()
    y = torch.randint(0, 10, (32,), device=dev, dtype=torch.int32)
    d = torch.zeros(32, 16, dtype=torch.bfloat16, device=dev)
    dt = torch.zeros(16, 32, dtype=torch.bfloat16, device=dev)
    db = torch.zeros(10, device=dev)
    loss = torch.zeros(1, device=dev)
    ops.softmax_xent(logits, 16, y, 32, 10, d, 16, dt, 32, db, loss)
    esum, ecorrect = torch.zeros(1, device=dev), torch.zeros(1, dtype=torch.int32, device=dev)
    ops.eval_reduce(logits, 16, y, 32, 10, esum, ecorrect)
    torch.cuda.synchronize()
    oloss, od = oracle.softmax_xent(logits[:, :10], y)
    assert abs(float(loss) - oloss) < 1e-3 * max(1.0, abs(oloss))
    assert torch.allclose(d[:, :10].float().cpu(), od, atol=2e-3)
    el, ec = oracle.evaluate(logits[:, :10], y)
    assert abs(float(esum) - el) < 1e-3 * max(1.0, abs(el)) and int(ecorrect) == ec
