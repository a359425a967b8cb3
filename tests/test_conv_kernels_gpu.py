"""Convolution-path kernels (im2col / col2im / pooling / BatchNorm / avg-pool) against PyTorch fp32."""

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ops():
    from p2pfl_b200 import ops as _ops

    _ops.lib()
    return _ops


def _nhwc(x):  # NCHW fp32 -> NHWC bf16
    return x.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)


@pytest.mark.parametrize("B,C,H,k,stride,pad,Cout", [(4, 1, 28, 5, 1, 2, 32), (3, 32, 14, 5, 1, 2, 64), (2, 16, 8, 3, 2, 1, 24), (2, 8, 8, 1, 2, 0, 16)])
def test_conv_as_im2col_gemm_fwd_wgrad_dgrad(ops, B, C, H, k, stride, pad, Cout):
    torch.manual_seed(0)
    x = torch.randn(B, C, H, H, device=DEV).to(torch.bfloat16).float()
    w = (torch.randn(Cout, C, k, k, device=DEV) / (C * k * k) ** 0.5).to(torch.bfloat16).float()
    bias = torch.randn(Cout, device=DEV)
    Ho = ops.conv_out(H, k, stride, pad)
    pix, K = B * Ho * Ho, C * k * k
    ldk, ldp = (K + 7) // 8 * 8, (pix + 7) // 8 * 8
    xn = _nhwc(x)
    col = torch.zeros(pix, ldk, device=DEV, dtype=torch.bfloat16)
    colT = torch.zeros(K, ldp, device=DEV, dtype=torch.bfloat16)
    ops.im2col(xn, B, H, H, C, k, k, stride, pad, col, ldk, colT, ldp)
    ref_col = F.unfold(x, k, padding=pad, stride=stride).transpose(1, 2).reshape(pix, K)
    assert torch.equal(col[:, :K].float(), ref_col)
    assert torch.equal(colT[:, :pix], col[:, :K].t())
    # forward: out[pix, Cout] = col . Wmat^T + bias, ReLU
    wmat = torch.zeros(Cout, ldk, device=DEV, dtype=torch.bfloat16)
    ops.pad_cast(w.reshape(Cout, K).contiguous(), Cout, K, wmat, ldk)
    out = torch.zeros(pix, Cout, device=DEV, dtype=torch.bfloat16)
    ops.gemm_tn(col, ldk, wmat, ldk, pix, Cout, K, bias=bias, bias_per_row=False, relu=True, out_bf16=out, ld_bf16=Cout)
    ref = F.relu(F.conv2d(x, w, bias, stride=stride, padding=pad))
    assert torch.allclose(out.float().view(B, Ho, Ho, Cout).permute(0, 3, 1, 2), ref, atol=3e-2, rtol=3e-2)
    # weight gradient: dW[Cout, K] = dOut^T . col   (K dimension = pixels)
    dout = torch.randn(pix, Cout, device=DEV).to(torch.bfloat16)
    doutT = torch.zeros(Cout, ldp, device=DEV, dtype=torch.bfloat16)
    doutT[:, :pix] = dout.t()
    dw = torch.zeros(Cout, K, device=DEV)
    ops.gemm_tn(doutT, ldp, colT, ldp, Cout, K, pix, out_f32=dw, ld_f32=K)
    xr = x.clone().requires_grad_(True)
    wr = w.clone().requires_grad_(True)
    y = F.conv2d(xr, wr, None, stride=stride, padding=pad)
    y.backward(dout.float().view(B, Ho, Ho, Cout).permute(0, 3, 1, 2))
    rel = (dw.view_as(wr.grad) - wr.grad).norm() / wr.grad.norm()
    assert rel < 2e-2, float(rel)
    # data gradient: dcol = dOut . Wmat ; dx = col2im(dcol)
    ldc = (Cout + 7) // 8 * 8
    wT = torch.zeros(K, ldc, device=DEV, dtype=torch.bfloat16)
    wT[:, :Cout] = w.reshape(Cout, K).t().to(torch.bfloat16)
    doutp = torch.zeros(pix, ldc, device=DEV, dtype=torch.bfloat16)
    doutp[:, :Cout] = dout
    dcol = torch.zeros(pix, ldk, device=DEV, dtype=torch.bfloat16)
    ops.gemm_tn(doutp, ldc, wT, ldc, pix, K, Cout, out_bf16=dcol, ld_bf16=ldk)
    dx = torch.zeros(B, H, H, C, device=DEV, dtype=torch.bfloat16)
    ops.col2im(dcol, ldk, B, H, H, C, k, k, stride, pad, dx)
    rel = (dx.float().permute(0, 3, 1, 2) - xr.grad).norm() / xr.grad.norm()
    assert rel < 3e-2, float(rel)


def test_maxpool_fwd_and_relu_bwd(ops):
    torch.manual_seed(1)
    B, C, H = 3, 16, 8
    pre = torch.randn(B, C, H, H, device=DEV).to(torch.bfloat16).float()
    act = F.relu(pre)
    an = _nhwc(act)
    Ho = H // 2
    y = torch.zeros(B, Ho, Ho, C, device=DEV, dtype=torch.bfloat16)
    arg = torch.zeros(B, Ho, Ho, C, device=DEV, dtype=torch.uint8)
    flat = torch.zeros(B, C * Ho * Ho, device=DEV, dtype=torch.bfloat16)
    flatT = torch.zeros(C * Ho * Ho, 8, device=DEV, dtype=torch.bfloat16)
    ops.maxpool2_fwd(an, B, H, H, C, y, arg, flat, flatT, 8)
    ref = F.max_pool2d(act, 2, 2)
    assert torch.equal(y.float().permute(0, 3, 1, 2), ref)
    assert torch.equal(flat.float(), ref.reshape(B, -1))
    assert torch.equal(flatT[:, :B], flat.t())
    # backward through pool and ReLU, gradient arriving in NCHW-flatten order
    dy = torch.randn(B, C * Ho * Ho, device=DEV).to(torch.bfloat16)
    pix = B * H * H
    dx = torch.zeros(pix, C, device=DEV, dtype=torch.bfloat16)
    dxT = torch.zeros(C, pix, device=DEV, dtype=torch.bfloat16)
    db = torch.zeros(C, device=DEV)
    ops.maxpool2_relu_bwd(None, dy, arg, an, B, H, H, C, dx, dxT, pix, db)
    pr = pre.clone().requires_grad_(True)
    F.max_pool2d(F.relu(pr), 2, 2).backward(dy.float().view(B, C, Ho, Ho))
    got = dx.float().view(B, H, H, C).permute(0, 3, 1, 2)
    # ties (several equal maxima, e.g. zeros after ReLU) may pick another winner: compare where unambiguous
    assert torch.allclose(got.sum((2, 3)), pr.grad.sum((2, 3)), atol=0.3, rtol=5e-2)
    assert (got != pr.grad).float().mean() < 0.02
    assert torch.equal(dxT, dx.t())
    assert torch.allclose(db, dx.float().sum(0), atol=1e-3)
    # NHWC-shaped incoming gradient takes the other entry
    dy2 = dy.view(B, C, Ho, Ho).permute(0, 2, 3, 1).contiguous()
    dx2 = torch.zeros_like(dx)
    ops.maxpool2_relu_bwd(dy2, None, arg, an, B, H, H, C, dx2, None, 0, None)
    assert torch.equal(dx2, dx)


@pytest.mark.parametrize("relu,res", [(False, False), (True, False), (True, True)])
def test_batchnorm_fwd_bwd(ops, relu, res):
    torch.manual_seed(2)
    B, C, H = 8, 24, 6
    x = torch.randn(B, C, H, H, device=DEV).to(torch.bfloat16).float() * 2 + 0.5
    r = torch.randn(B, C, H, H, device=DEV).to(torch.bfloat16).float() if res else None
    bn = torch.nn.BatchNorm2d(C).to(DEV)
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
    rm, rv = bn.running_mean.clone(), bn.running_var.clone()
    xr = x.clone().requires_grad_(True)
    rr = r.clone().requires_grad_(True) if res else None
    out = bn(xr)
    if res:
        out = out + rr
    if relu:
        out = F.relu(out)
    dy = torch.randn_like(out).to(torch.bfloat16).float()
    out.backward(dy)
    pixels = B * H * H
    y = torch.zeros(B, H, H, C, device=DEV, dtype=torch.bfloat16)
    sm, sr = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    ops.bn_fwd(_nhwc(x), pixels, C, bn.weight.detach(), bn.bias.detach(), rm, rv, 0.1, 1e-5, relu, _nhwc(r) if res else None, y, sm, sr)
    assert torch.allclose(y.float().permute(0, 3, 1, 2), out.detach(), atol=3e-2, rtol=3e-2)
    assert torch.allclose(rm, bn.running_mean, atol=1e-4) and torch.allclose(rv, bn.running_var, atol=1e-3)
    dx = torch.zeros_like(y)
    dres = torch.zeros_like(y) if res else None
    dg, dbt = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    dxT = torch.zeros(C, pixels, device=DEV, dtype=torch.bfloat16)
    ops.bn_bwd(_nhwc(x), y, _nhwc(dy), pixels, C, bn.weight.detach(), sm, sr, relu, dx, dres, dg, dbt, dxT, pixels)
    assert torch.equal(dxT, dx.view(pixels, C).t())
    # the ReLU mask is taken from the bf16 output: elements within rounding of zero may flip
    rel = (dx.float().permute(0, 3, 1, 2) - xr.grad).norm() / xr.grad.norm()
    assert rel < 5e-2, float(rel)
    assert (dg - bn.weight.grad).norm() / bn.weight.grad.norm() < 5e-2
    assert (dbt - bn.bias.grad).norm() / bn.bias.grad.norm() < 5e-2
    if res:
        assert (dres.float().permute(0, 3, 1, 2) - rr.grad).norm() / rr.grad.norm() < 5e-2


def test_batchnorm_inference(ops):
    B, C, H = 4, 40, 5
    x = torch.randn(B, C, H, H, device=DEV).to(torch.bfloat16).float()
    bn = torch.nn.BatchNorm2d(C).to(DEV).eval()
    with torch.no_grad():
        bn.running_mean.uniform_(-1, 1)
        bn.running_var.uniform_(0.5, 2)
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
        ref = F.relu(bn(x))
    y = torch.zeros(B, H, H, C, device=DEV, dtype=torch.bfloat16)
    ops.bn_infer(_nhwc(x), B * H * H, C, bn.weight.detach(), bn.bias.detach(), bn.running_mean, bn.running_var, 1e-5, True, None, y)
    assert torch.allclose(y.float().permute(0, 3, 1, 2), ref, atol=3e-2, rtol=3e-2)


def test_col2im_adds_the_residual_gradient(ops):
    B, C, H, k = 2, 8, 6, 3
    K = C * k * k
    dcol = torch.randn(B * H * H, K, device=DEV).to(torch.bfloat16)
    extra = torch.randn(B, H, H, C, device=DEV).to(torch.bfloat16)
    a = torch.zeros(B, H, H, C, device=DEV, dtype=torch.bfloat16)
    b = torch.zeros_like(a)
    ops.col2im(dcol, K, B, H, H, C, k, k, 1, 1, a)
    ops.col2im(dcol, K, B, H, H, C, k, k, 1, 1, b, extra)
    assert torch.allclose(b.float(), a.float() + extra.float(), atol=3e-2, rtol=2e-2)


def test_global_avgpool(ops):
    B, C, HW = 5, 48, 16
    x = torch.randn(B, HW, C, device=DEV).to(torch.bfloat16)
    y = torch.zeros(B, C, device=DEV, dtype=torch.bfloat16)
    yT = torch.zeros(C, 8, device=DEV, dtype=torch.bfloat16)
    ops.avgpool_fwd(x, B, HW, C, y, yT, 8)
    assert torch.allclose(y.float(), x.float().mean(1), atol=1e-2)
    assert torch.equal(yT[:, :B], y.t())
    dy = torch.randn(B, C, device=DEV).to(torch.bfloat16)
    dx = torch.zeros(B, HW, C, device=DEV, dtype=torch.bfloat16)
    ops.avgpool_bwd(dy, B, HW, C, dx)
    assert torch.allclose(dx.float(), (dy.float() / HW)[:, None, :].expand(B, HW, C), atol=1e-3)


def test_nchw_to_nhwc(ops):
    x = torch.randn(3, 5, 4, 6, device=DEV)
    y = torch.zeros(3, 4, 6, 5, device=DEV, dtype=torch.bfloat16)
    ops.nchw_to_nhwc(x, 3, 5, 4, 6, y)
    assert torch.equal(y, _nhwc(x))


# ------------------------------------------------------- tap-major ("kc") path
@pytest.mark.parametrize("B,C,H,k,stride,pad,Cout", [(4, 64, 16, 3, 1, 1, 64), (3, 64, 16, 3, 2, 1, 128), (2, 128, 8, 1, 2, 0, 256), (5, 8, 6, 3, 1, 1, 16)])
def test_conv_tap_major_fwd_wgrad_dgrad(ops, B, C, H, k, stride, pad, Cout):
    """The ResNet conv path: vectorised im2col, packed weights, MN-major backward GEMMs
    (no transposed copies), split-K wgrad, vectorised col2im (+ residual), unpacked gradient."""
    torch.manual_seed(1)
    x = torch.randn(B, C, H, H, device=DEV).to(torch.bfloat16).float()
    w = (torch.randn(Cout, C, k, k, device=DEV) / (C * k * k) ** 0.5).to(torch.bfloat16).float()
    Ho = ops.conv_out(H, k, stride, pad)
    pix, K, taps = B * Ho * Ho, C * k * k, k * k
    xn = _nhwc(x)
    col = torch.full((pix, K), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.im2col_kc(xn, B, H, H, C, k, k, stride, pad, col, K)
    ref_col = F.unfold(x, k, padding=pad, stride=stride).transpose(1, 2).reshape(pix, C, taps)
    assert torch.equal(col.float().view(pix, taps, C), ref_col.permute(0, 2, 1))
    # packed weights: flat fp32 (PyTorch order, at an offset) -> bf16 [Cout, K] tap-major
    off = 40
    flat = torch.zeros(off + Cout * K, device=DEV)
    flat[off:] = w.reshape(-1)
    table = torch.tensor([[off, 64, Cout, C, taps, K, 1, 0]], dtype=torch.int64, device=DEV)
    wk = torch.zeros(64 + Cout * K, device=DEV, dtype=torch.bfloat16)
    ops.conv_pack(flat, wk, table, 1, Cout * K)
    assert torch.equal(wk[64:].float().view(Cout, taps, C), w.view(Cout, C, taps).permute(0, 2, 1))
    wmat = wk[64:]
    out = torch.zeros(pix, Cout, device=DEV, dtype=torch.bfloat16)
    ops.gemm_tn(col, K, wmat, K, pix, Cout, K, out_bf16=out, ld_bf16=Cout)
    ref = F.conv2d(x, w, None, stride=stride, padding=pad)
    assert torch.allclose(out.float().view(B, Ho, Ho, Cout).permute(0, 3, 1, 2), ref, atol=3e-2, rtol=3e-2)
    # backward: both GEMMs in one launch, reading dout / col / wk in place
    dout = torch.randn(pix, Cout, device=DEV).to(torch.bfloat16)
    gw = torch.zeros(64 + Cout * K, device=DEV)
    dcol = torch.zeros(pix, K, device=DEV, dtype=torch.bfloat16)
    ops.gemm_tn_grouped(
        torch.device(DEV),
        [
            ops.gemm_args(dout, Cout, col, K, Cout, K, pix, out_f32=gw[64:], ld_f32=K, split_k=-1, a_mn=True, b_mn=True),
            ops.gemm_args(dout, Cout, wmat, K, pix, K, Cout, out_bf16=dcol, ld_bf16=K, b_mn=True),
        ],
    )
    grads = torch.zeros_like(flat)
    ops.conv_unpack(gw, grads, table, 1, Cout * K)
    extra = torch.randn(B, H, H, C, device=DEV).to(torch.bfloat16)
    dx = torch.zeros(B, H, H, C, device=DEV, dtype=torch.bfloat16)
    ops.col2im_kc(dcol, K, B, H, H, C, k, k, stride, pad, dx, extra)
    xr = x.clone().requires_grad_(True)
    wr = w.clone().requires_grad_(True)
    y = F.conv2d(xr, wr, None, stride=stride, padding=pad)
    y.backward(dout.float().view(B, Ho, Ho, Cout).permute(0, 3, 1, 2))
    rel = (grads[off:].view_as(wr.grad) - wr.grad).norm() / wr.grad.norm()
    assert rel < 2e-2, float(rel)
    assert float(grads[:off].abs().max()) == 0.0
    ref_dx = xr.grad.permute(0, 2, 3, 1) + extra.float()
    rel = (dx.float() - ref_dx).norm() / ref_dx.norm()
    assert rel < 2e-2, float(rel)


def test_conv_pack_keeps_pytorch_order_for_the_stem(ops):
    """tap_major = 0 rows (3 input channels) are a plain padded cast; several rows per launch."""
    torch.manual_seed(2)
    w0 = torch.randn(64, 3, 3, 3, device=DEV)
    w1 = torch.randn(16, 8, 3, 3, device=DEV)
    flat = torch.cat([w0.reshape(-1), w1.reshape(-1)])
    table = torch.tensor([[0, 0, 64, 3, 9, 32, 0, 0], [w0.numel(), 64 * 32, 16, 8, 9, 72, 1, 0]], dtype=torch.int64, device=DEV)
    wk = torch.zeros(64 * 32 + 16 * 72, device=DEV, dtype=torch.bfloat16)
    ops.conv_pack(flat, wk, table, 2, max(w0.numel(), w1.numel()))
    a = wk[: 64 * 32].view(64, 32)
    assert torch.equal(a[:, :27], w0.reshape(64, 27).to(torch.bfloat16)) and float(a[:, 27:].abs().max()) == 0.0
    assert torch.equal(wk[64 * 32 :].view(16, 9, 8), w1.view(16, 8, 9).permute(0, 2, 1).to(torch.bfloat16))
    g = torch.zeros_like(flat)
    ops.conv_unpack(wk.float(), g, table, 2, max(w0.numel(), w1.numel()))
    assert torch.equal(g, flat.to(torch.bfloat16).float())
