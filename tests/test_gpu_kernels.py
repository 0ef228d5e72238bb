"""Numerics of every hand-written sm_100a kernel against a plain PyTorch fp32 reference of
the same op.  Activations are bf16 in the fused path, so the
comparisons use relative-L2 bands sized for bf16 rounding, not bit equality."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-12))


def report(name, err, tol):
    print(f"[numerics] {name}: rel-L2 {err:.3e} (tol {tol:.1e})")
    assert err < tol, f"{name}: rel-L2 {err:.3e} >= {tol:.1e}"


@pytest.fixture(scope="module")
def ops():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    from torch_distributed_sandbox_b200 import ops as o
    return o


def nhwc_bf16(x_nchw):
    return x_nchw.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)


def to_nchw(x_nhwc):
    return x_nhwc.float().permute(0, 3, 1, 2).contiguous()


def test_resize_matches_pil(ops):
    from torch_distributed_sandbox_b200.data.resize import pil_bilinear_table, resize_u8_reference
    rng = np.random.default_rng(0)
    src = rng.integers(0, 256, (3, 28, 28), dtype=np.uint8)
    for (H, W) in [(64, 96), (300, 500)]:
        xt = torch.from_numpy(pil_bilinear_table(28, W)).to(DEV)
        yt = torch.from_numpy(pil_bilinear_table(28, H)).to(DEV)
        dst = torch.empty(3, H, W, device=DEV)
        ops.resize_pil_bilinear(torch.from_numpy(src).to(DEV), xt, yt, dst)
        ref = torch.from_numpy(resize_u8_reference(src, (H, W))).float() / 255.0
        assert torch.equal(dst.cpu(), ref), f"resize {H}x{W} is not bit-exact"


@pytest.mark.parametrize("shape", [(2, 64, 96), (1, 36, 200), (3, 128, 132), (1, 300, 72), (2, 4, 8)])
@pytest.mark.parametrize("impl", ["tc", "ffma"])
def test_conv1_fwd(ops, shape, impl):
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(1)
    x = torch.rand(B, H, W, device=DEV, generator=g)
    w = torch.randn(16, 1, 5, 5, device=DEV, generator=g) * 0.2
    b = torch.randn(16, device=DEV, generator=g) * 0.1
    x1 = torch.empty(B, H, W, 16, device=DEV, dtype=torch.bfloat16)
    part = torch.zeros(ops.conv1_fwd_grid(), 32, device=DEV)
    x1.fill_(float("nan"))
    if impl == "tc":
        n = ops.conv1_fwd_tc(x, w, b, x1, part)
    else:
        ops.conv1_fwd(x, w, b, x1, part)
        n = ops.conv1_fwd_grid()
    ref = F.conv2d(x[:, None], w, b, padding=2)
    report(f"conv1_fwd[{impl}]", rel(to_nchw(x1), ref), 4e-3)
    s = part[:n].double().sum(0)
    r = x1.float().double()          # statistics are those of the stored (bf16-rounded) tensor
    report(f"conv1_fwd[{impl}] sum", rel(s[:16], r.sum((0, 1, 2))), 1e-5)
    report(f"conv1_fwd[{impl}] sumsq", rel(s[16:], (r ** 2).sum((0, 1, 2))), 1e-5)


def test_bn_finalize(ops):
    C, n = 32, 7
    g = torch.Generator(device=DEV).manual_seed(2)
    vals = torch.randn(n, 1000, C, device=DEV, generator=g) * 2 + 0.5
    part = torch.cat([vals.sum(1), (vals * vals).sum(1)], dim=1).contiguous()
    gamma = torch.rand(C, device=DEV, generator=g) + 0.5
    beta = torch.randn(C, device=DEV, generator=g)
    rm, rv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    nbt = torch.zeros((), device=DEV, dtype=torch.int64)
    out = torch.zeros(4, C, device=DEV)
    ops.bn_finalize(part, n, C, n * 1000, gamma, beta, rm, rv, nbt, out)
    flat = vals.reshape(-1, C).double()
    mean, var = flat.mean(0), flat.var(0, unbiased=False)
    rstd = 1 / torch.sqrt(var + 1e-5)
    report("bn scale", rel(out[0], gamma.double() * rstd), 1e-5)
    report("bn shift", rel(out[1], beta.double() - mean * gamma.double() * rstd), 1e-4)
    report("bn mean", rel(out[2], mean), 1e-5)
    report("bn running_mean", rel(rm, 0.1 * mean), 1e-5)
    report("bn running_var", rel(rv, 0.9 + 0.1 * flat.var(0, unbiased=True)), 1e-5)
    assert int(nbt) == 1


@pytest.mark.parametrize("C", [16, 32])
def test_bn_relu_pool(ops, C):
    B, H, W = 2, 20, 36
    g = torch.Generator(device=DEV).manual_seed(3)
    x = torch.randn(B, C, H, W, device=DEV, generator=g)
    xb = nhwc_bf16(x)
    bn = torch.zeros(4, C, device=DEV)
    bn[0] = torch.randn(C, device=DEV, generator=g)          # scale (both signs)
    bn[1] = torch.randn(C, device=DEV, generator=g) * 0.3     # shift
    p = torch.empty(B, H // 2, W // 2, C, device=DEV, dtype=torch.bfloat16)
    ops.bn_relu_pool(xb, bn, p)
    y = to_nchw(xb) * bn[0].view(1, C, 1, 1) + bn[1].view(1, C, 1, 1)
    ref = F.max_pool2d(F.relu(y), 2)
    report(f"bn_relu_pool C={C}", rel(to_nchw(p), ref), 4e-3)


@pytest.fixture(params=[5, 4, 3, 2, 1], ids=["rowstat_tma_swizzled_rows", "rowstat_cpasync_wgradM128", "rowstat_cpasync",
                                              "rowstat_tma", "tapstat_tma"])
def tc_version(ops, request):
    ops.conv_tc_set_version(request.param)
    yield request.param
    ops.conv_tc_set_version(5)


@pytest.mark.parametrize("shape", [(2, 24, 200), (1, 50, 128), (3, 17, 300), (1, 3, 40), (2, 133, 260)])
def test_conv2_fwd_tc(ops, shape, tc_version):
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(4)
    x = torch.randn(B, 16, H, W, device=DEV, generator=g)
    w = torch.randn(32, 16, 5, 5, device=DEV, generator=g) * 0.05
    b = torch.randn(32, device=DEV, generator=g) * 0.1
    xb = nhwc_bf16(x)
    out = torch.full((B, H, W, 32), float("nan"), device=DEV, dtype=torch.bfloat16)
    part = torch.zeros(ops.conv_tc_max_grid(), 64, device=DEV)
    n = ops.conv2_fwd_tc(xb, w, b, out, part)
    torch.cuda.synchronize()
    ref = F.conv2d(to_nchw(xb), w.to(torch.bfloat16).float(), b, padding=2)
    report(f"conv2_fwd_tc {shape}", rel(to_nchw(out), ref), 5e-3)
    r = out.float().double()
    s = part[:n].double().sum(0)
    report("conv2_fwd_tc sum", rel(s[:32], r.sum((0, 1, 2))), 1e-4)
    report("conv2_fwd_tc sumsq", rel(s[32:], (r ** 2).sum((0, 1, 2))), 1e-4)


@pytest.fixture
def few_ctas(ops):
    """Leave the persistent tcgen05 kernels 2 SMs: every CTA then walks several long work items back to back, so
    the accumulator / operand rings wrap many times and carry their position across items - the situation of the
    full-size step, which the small shapes on 148 SMs (one short item per CTA) never reach."""
    ops.set_reserved_sms(ops.sm_count() - 2)
    yield
    ops.set_reserved_sms(0)


@pytest.mark.parametrize("ver", [5, 4, 3, 2])
def test_conv2_tc_long_items(ops, few_ctas, ver):
    ops.conv_tc_set_version(ver)
    try:
        B, H, W = 3, 150, 300
        g = torch.Generator(device=DEV).manual_seed(21)
        x = torch.randn(B, 16, H, W, device=DEV, generator=g)
        w = torch.randn(32, 16, 5, 5, device=DEV, generator=g) * 0.05
        b = torch.randn(32, device=DEV, generator=g) * 0.1
        xb = nhwc_bf16(x)
        out = torch.full((B, H, W, 32), float("nan"), device=DEV, dtype=torch.bfloat16)
        part = torch.zeros(ops.conv_tc_max_grid(), 64, device=DEV)
        n = ops.conv2_fwd_tc(xb, w, b, out, part)
        torch.cuda.synchronize()
        assert n <= 2
        ref = F.conv2d(to_nchw(xb), w.to(torch.bfloat16).float(), b, padding=2)
        report(f"conv2_fwd_tc long items v{ver}", rel(to_nchw(out), ref), 5e-3)
        s = part[:n].double().sum(0)
        report("conv2_fwd_tc long items sum", rel(s[:32], out.float().double().sum((0, 1, 2))), 1e-4)
        dy = nhwc_bf16(torch.randn(B, 32, H, W, device=DEV, generator=g))
        dx = torch.full((B, H, W, 16), float("nan"), device=DEV, dtype=torch.bfloat16)
        ops.conv2_dgrad_tc(dy, w, dx)
        torch.cuda.synchronize()
        refd = torch.nn.grad.conv2d_input((B, 16, H, W), w.to(torch.bfloat16).float(), to_nchw(dy), padding=2)
        report(f"conv2_dgrad_tc long items v{ver}", rel(to_nchw(dx), refd), 5e-3)
        pw = torch.zeros(ops.conv_tc_max_grid(), 32 * 16 * 25, device=DEV)
        dw = torch.zeros(32, 16, 5, 5, device=DEV)
        ops.conv2_wgrad_tc(xb, dy, pw, dw)
        torch.cuda.synchronize()
        refw = torch.nn.grad.conv2d_weight(to_nchw(xb), (32, 16, 5, 5), to_nchw(dy), padding=2)
        report(f"conv2_wgrad_tc long items v{ver}", rel(dw, refw), 5e-3)
    finally:
        ops.conv_tc_set_version(5)


def test_conv1_tc_many_items_per_cta(ops, few_ctas):
    """conv1 forward on 2 CTAs: each walks ~10 tiles, so the input-stage and TMEM rings carry across items."""
    B, H, W = 2, 150, 300
    g = torch.Generator(device=DEV).manual_seed(22)
    x = torch.rand(B, H, W, device=DEV, generator=g)
    w = torch.randn(16, 1, 5, 5, device=DEV, generator=g) * 0.2
    b = torch.randn(16, device=DEV, generator=g) * 0.1
    x1 = torch.full((B, H, W, 16), float("nan"), device=DEV, dtype=torch.bfloat16)
    part = torch.zeros(ops.conv1_fwd_grid(), 32, device=DEV)
    n = ops.conv1_fwd_tc(x, w, b, x1, part)
    torch.cuda.synchronize()
    assert n <= 2
    report("conv1_fwd[tc] many items", rel(to_nchw(x1), F.conv2d(x[:, None], w, b, padding=2)), 4e-3)
    s = part[:n].double().sum(0)
    report("conv1_fwd[tc] many items sum", rel(s[:16], x1.float().double().sum((0, 1, 2))), 1e-5)


@pytest.mark.parametrize("shape", [(2, 24, 200), (1, 50, 128), (1, 2, 24), (2, 133, 260)])
def test_conv2_dgrad_tc(ops, shape, tc_version):
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(5)
    dy = torch.randn(B, 32, H, W, device=DEV, generator=g)
    w = torch.randn(32, 16, 5, 5, device=DEV, generator=g) * 0.05
    dyb = nhwc_bf16(dy)
    out = torch.full((B, H, W, 16), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.conv2_dgrad_tc(dyb, w, out)
    torch.cuda.synchronize()
    ref = torch.nn.grad.conv2d_input((B, 16, H, W), w.to(torch.bfloat16).float(), to_nchw(dyb), padding=2)
    report(f"conv2_dgrad_tc {shape}", rel(to_nchw(out), ref), 5e-3)


@pytest.mark.parametrize("shape", [(2, 24, 200), (1, 50, 128), (3, 17, 300), (1, 2, 24), (2, 133, 260)])
def test_conv2_wgrad_tc(ops, shape, tc_version):
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(6)
    x = torch.randn(B, 16, H, W, device=DEV, generator=g)
    dy = torch.randn(B, 32, H, W, device=DEV, generator=g)
    xb, dyb = nhwc_bf16(x), nhwc_bf16(dy)
    part = torch.zeros(ops.conv_tc_max_grid(), 32 * 16 * 25, device=DEV)
    dw = torch.zeros(32, 16, 5, 5, device=DEV)
    ops.conv2_wgrad_tc(xb, dyb, part, dw)
    torch.cuda.synchronize()
    ref = torch.nn.grad.conv2d_weight(to_nchw(xb), (32, 16, 5, 5), to_nchw(dyb), padding=2)
    report(f"conv2_wgrad_tc {shape}", rel(dw, ref), 2e-3)


@pytest.mark.parametrize("B", [1, 5, 8])
def test_fc_and_ce(ops, B):
    Fdim = 32 * 12 * 20
    g = torch.Generator(device=DEV).manual_seed(7)
    p2 = (torch.rand(B, Fdim, device=DEV, generator=g)).to(torch.bfloat16)
    W = torch.randn(10, Fdim, device=DEV, generator=g) * 0.02
    bias = torch.randn(10, device=DEV, generator=g) * 0.1
    labels = torch.randint(0, 10, (B,), device=DEV, generator=g)
    part = torch.zeros(ops.fc_grid(), 80, device=DEV)
    logits = torch.zeros(B, 10, device=DEV)
    ops.fc_fwd(p2, W, bias, part, logits)
    ref_logits = (p2.float() @ W.t() + bias).detach().requires_grad_(True)
    report(f"fc_fwd B={B}", rel(logits, ref_logits), 1e-5)
    loss, dl, db = torch.zeros((), device=DEV), torch.zeros(B, 10, device=DEV), torch.zeros(10, device=DEV)
    ops.cross_entropy(logits, labels, loss, dl, db)
    ref_loss = F.cross_entropy(ref_logits, labels)
    ref_loss.backward()
    report("ce loss", rel(loss, ref_loss.detach()), 1e-5)
    report("ce dlogits", rel(dl, ref_logits.grad), 1e-4)
    report("ce dbias", rel(db, ref_logits.grad.sum(0)), 1e-4)
    # backward: dW, dp2, and the fused-SGD variant
    dW = torch.zeros_like(W)
    dp2 = torch.empty_like(p2)
    W0 = W.clone()
    ops.fc_bwd(dl, p2, W, dW, dp2)
    report("fc_bwd dW", rel(dW, dl.t() @ p2.float()), 1e-5)
    report("fc_bwd dp2", rel(dp2, dl @ W0), 5e-3)
    assert torch.equal(W, W0)
    ops.fc_bwd(dl, p2, W, None, dp2, lr=0.5)
    report("fc_bwd fused sgd", rel(W, W0 - 0.5 * (dl.t() @ p2.float())), 1e-6)


@pytest.mark.parametrize("C", [16, 32])
def test_pool_bn_bwd(ops, C):
    B, H, W = 2, 24, 40
    g = torch.Generator(device=DEV).manual_seed(8)
    xb = nhwc_bf16(torch.randn(B, C, H, W, device=DEV, generator=g) * 1.5 + 0.3)
    dpb = nhwc_bf16(torch.randn(B, C, H // 2, W // 2, device=DEV, generator=g))
    gamma = torch.rand(C, device=DEV, generator=g) + 0.5
    gamma[::3] *= -1
    beta = torch.randn(C, device=DEV, generator=g) * 0.2
    # reference through autograd on the stored (bf16-rounded) tensor
    x = to_nchw(xb).requires_grad_(True)
    gam, bet = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    y = F.batch_norm(x, None, None, gam, bet, training=True, eps=1e-5)
    p = F.max_pool2d(F.relu(y), 2)
    p.backward(to_nchw(dpb))
    mean = x.detach().mean((0, 2, 3))
    var = x.detach().var((0, 2, 3), unbiased=False)
    rstd = 1 / torch.sqrt(var + 1e-5)
    bn = torch.stack([gamma * rstd, beta - mean * gamma * rstd, mean, rstd]).contiguous()
    grid = ops.pool_bn_bwd_grid()
    part = torch.zeros(grid, 64, device=DEV)
    red = torch.zeros(2, C, device=DEV)
    dg, db = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    ops.pool_bn_bwd(xb, dpb, bn, red, part, None, mode=0)
    ops.bn_bwd_finalize(part, grid, C, B * H * W, red, dg, db)
    report(f"bn bwd dgamma C={C}", rel(dg, gam.grad), 1e-3)
    report(f"bn bwd dbeta C={C}", rel(db, bet.grad), 1e-3)
    dx = torch.empty_like(xb)
    ops.pool_bn_bwd(xb, dpb, bn, red, None, dx, mode=1)
    report(f"pool_bn_bwd dx C={C}", rel(to_nchw(dx), x.grad), 6e-3)


@pytest.mark.parametrize("impl,shape", [("ffma", (2, 40, 72)), ("tc", (2, 40, 72)), ("tc", (1, 8, 128)),
                                        ("tc", (3, 36, 300)), ("tc", (2, 132, 260))])
def test_conv1_wgrad(ops, impl, shape):
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(9)
    xin = torch.rand(B, H, W, device=DEV, generator=g)
    w = torch.randn(16, 1, 5, 5, device=DEV, generator=g) * 0.2
    b = torch.zeros(16, device=DEV)
    gamma = torch.rand(16, device=DEV, generator=g) + 0.5
    beta = torch.randn(16, device=DEV, generator=g) * 0.2
    dpb = nhwc_bf16(torch.randn(B, 16, H // 2, W // 2, device=DEV, generator=g))
    x1 = torch.empty(B, H, W, 16, device=DEV, dtype=torch.bfloat16)
    part = torch.zeros(ops.conv1_fwd_grid(), 32, device=DEV)
    ops.conv1_fwd(xin, w, b, x1, part)
    bn = torch.zeros(4, 16, device=DEV)
    rm, rv = torch.zeros(16, device=DEV), torch.ones(16, device=DEV)
    ops.bn_finalize(part, ops.conv1_fwd_grid(), 16, B * H * W, gamma, beta, rm, rv, None, bn)
    grid = ops.pool_bn_bwd_grid()
    partb = torch.zeros(grid, 64, device=DEV)
    red = torch.zeros(2, 16, device=DEV)
    dg, db = torch.zeros(16, device=DEV), torch.zeros(16, device=DEV)
    ops.pool_bn_bwd(x1, dpb, bn, red, partb, None, mode=0)
    ops.bn_bwd_finalize(partb, grid, 16, B * H * W, red, dg, db)
    pw = torch.zeros(ops.conv1_wgrad_grid(), 400, device=DEV)
    dw = torch.zeros(16, 1, 5, 5, device=DEV)
    (ops.conv1_wgrad_tc if impl == "tc" else ops.conv1_wgrad)(xin, x1, dpb, bn, red, pw, dw)
    # Oracle on the SAME stored (bf16-rounded) conv output the kernel sees: otherwise the few
    # windows whose arg-max flips under bf16 rounding dominate the error of a noise-like sum.
    x1r = to_nchw(x1).requires_grad_(True)
    y = F.batch_norm(x1r, None, None, gamma, beta, training=True)
    F.max_pool2d(F.relu(y), 2).backward(to_nchw(dpb))
    ref = torch.nn.grad.conv2d_weight(xin[:, None], (16, 1, 5, 5), x1r.grad, padding=2)
    # the tcgen05 path rounds the conv1-output gradient and the input to bf16 (fp32 accumulate)
    report(f"conv1_wgrad[{impl}] {shape}", rel(dw, ref), 2e-3 if impl == "ffma" else 8e-3)


def test_sgd_and_permute(ops):
    g = torch.Generator(device=DEV).manual_seed(10)
    p = torch.randn(13354, device=DEV, generator=g)
    gr = torch.randn(13354, device=DEV, generator=g)
    ref = p - 0.25 * gr
    ops.sgd(p, gr, 0.25)
    report("sgd", rel(p, ref), 1e-7)
    C, Hd, Wd = 32, 6, 10
    src = torch.randn(10, C * Hd * Wd, device=DEV, generator=g)
    hwc, back = torch.empty_like(src), torch.empty_like(src)
    ops.fcw_permute(src, hwc, C, Hd, Wd, True)
    ref_hwc = src.view(10, C, Hd, Wd).permute(0, 2, 3, 1).reshape(10, -1)
    assert torch.equal(hwc, ref_hwc)
    ops.fcw_permute(hwc, back, C, Hd, Wd, False)
    assert torch.equal(back, src)


def _bn_case(ops, C, B, H, W, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    xb = nhwc_bf16(torch.randn(B, C, H, W, device=DEV, generator=g) * 1.3 + 0.2)
    gamma = torch.rand(C, device=DEV, generator=g) + 0.5
    gamma[1::4] *= -1
    beta = torch.randn(C, device=DEV, generator=g) * 0.3
    xf = xb.float()
    mean, var = xf.mean((0, 1, 2)), xf.var((0, 1, 2), unbiased=False)
    rstd = 1 / torch.sqrt(var + 1e-5)
    bn = torch.stack([gamma * rstd, beta - mean * gamma * rstd, mean, rstd]).contiguous()
    p = torch.empty(B, H // 2, W // 2, C, device=DEV, dtype=torch.bfloat16)
    ops.bn_relu_pool(xb, bn, p)
    return g, xb, bn, p


def test_fc_bwd_fused_bn2_backward_stats(ops):
    """sum(dy), sum(dy*xhat) produced inside fc_bwd from (dp2, p2) == the separate pass over x2."""
    B, H, W, C = 5, 16, 24, 32
    g, xb, bn, p2 = _bn_case(ops, C, B, H, W, 21)
    Fd = C * (H // 2) * (W // 2)
    Wt = torch.randn(10, Fd, device=DEV, generator=g) * 0.05
    dl = torch.randn(B, 10, device=DEV, generator=g) * 0.1
    dp2 = torch.empty_like(p2)
    part = torch.zeros(ops.fc_grid(), 64, device=DEV)
    ops.fc_bwd(dl, p2.view(B, Fd), Wt, None, dp2.view(B, Fd), bn2=bn, bwd_partials=part)
    red_f, dg_f, db_f = torch.zeros(2, C, device=DEV), torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    ops.bn_bwd_finalize(part, ops.fc_grid(), C, B * H * W, red_f, dg_f, db_f)
    grid = ops.pool_bn_bwd_grid()
    part2 = torch.zeros(grid, 64, device=DEV)
    red_r, dg_r, db_r = torch.zeros(2, C, device=DEV), torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    ops.pool_bn_bwd(xb, dp2, bn, red_r, part2, None, mode=0)
    ops.bn_bwd_finalize(part2, grid, C, B * H * W, red_r, dg_r, db_r)
    report("fused bn2 bwd dbeta", rel(db_f, db_r), 1e-5)
    report("fused bn2 bwd dgamma", rel(dg_f, dg_r), 1e-2)


@pytest.mark.parametrize("shape", [(2, 24, 200), (1, 50, 128)])
def test_conv2_dgrad_fused_bn1_backward_stats(ops, shape):
    """dgrad epilogue reductions (dp1 with p1 read-before-overwrite, in place) == the separate pass over x1."""
    B, H, W = shape                                  # pooled-resolution geometry of layer 1
    g, x1, bn1, p1 = _bn_case(ops, 16, B, 2 * H, 2 * W, 22)
    dy = nhwc_bf16(torch.randn(B, 32, H, W, device=DEV, generator=g))
    w = torch.randn(32, 16, 5, 5, device=DEV, generator=g) * 0.05
    ref_dp1 = torch.empty(B, H, W, 16, device=DEV, dtype=torch.bfloat16)
    ops.conv2_dgrad_tc(dy, w, ref_dp1)
    part = torch.zeros(ops.conv_tc_max_grid(), 32, device=DEV)
    inplace = p1.clone()
    n = ops.conv2_dgrad_stats_tc(dy, w, inplace, bn1, inplace, part)     # p1 buffer becomes dp1
    assert torch.equal(inplace, ref_dp1)
    red_f, dg_f, db_f = torch.zeros(2, 16, device=DEV), torch.zeros(16, device=DEV), torch.zeros(16, device=DEV)
    ops.bn_bwd_finalize(part, n, 16, B * 4 * H * W, red_f, dg_f, db_f)
    grid = ops.pool_bn_bwd_grid()
    part2 = torch.zeros(grid, 64, device=DEV)
    red_r, dg_r, db_r = torch.zeros(2, 16, device=DEV), torch.zeros(16, device=DEV), torch.zeros(16, device=DEV)
    ops.pool_bn_bwd(x1, ref_dp1, bn1, red_r, part2, None, mode=0)
    ops.bn_bwd_finalize(part2, grid, 16, B * 4 * H * W, red_r, dg_r, db_r)
    report(f"fused bn1 bwd dbeta {shape}", rel(db_f, db_r), 1e-5)
    report(f"fused bn1 bwd dgamma {shape}", rel(dg_f, dg_r), 1e-2)


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("B", [9, 33])
def test_fc_bwd_any_batch_single_pass(ops, B, dt):
    """Batches > 8 run the any-batch kernel (one pass over W): dW, dp2, fused SGD and the fused BN-2 reductions
    against fp64 matmuls / the <= 8-row kernel."""
    C, H, W_ = 32, 8, 12
    Fdim = C * H * W_
    g = torch.Generator(device=DEV).manual_seed(17)
    p2 = torch.rand(B, Fdim, device=DEV, generator=g).to(dt)
    Wt = torch.randn(10, Fdim, device=DEV, generator=g) * 0.02
    dl = torch.randn(B, 10, device=DEV, generator=g) * 0.1
    bn = torch.stack([torch.rand(C, device=DEV, generator=g) + 0.5, torch.randn(C, device=DEV, generator=g) * 0.1,
                      torch.randn(C, device=DEV, generator=g) * 0.1, torch.rand(C, device=DEV, generator=g) + 0.5]).contiguous()
    dW, dp2, W0 = torch.zeros_like(Wt), torch.empty_like(p2), Wt.clone()
    part = torch.zeros(ops.fc_grid(), 64, device=DEV)
    ops.fc_bwd(dl, p2, Wt, dW, dp2, bn2=bn, bwd_partials=part)
    report(f"fc_bwd any-batch dW B={B}", rel(dW, dl.double().t() @ p2.double()), 1e-5)
    report(f"fc_bwd any-batch dp2 B={B}", rel(dp2, dl.double() @ W0.double()), 5e-3 if dt == torch.bfloat16 else 1e-6)
    # BN-2 reductions: the same rows through the <= 8-row kernel, chunk by chunk
    ref_part = torch.zeros_like(part)
    tmp = torch.empty_like(p2)
    for b0 in range(0, B, 8):
        ops.fc_bwd(dl[b0:b0 + 8], p2[b0:b0 + 8], Wt, None, tmp[b0:b0 + 8], accumulate=b0 > 0, bn2=bn, bwd_partials=ref_part)
    assert torch.equal(tmp, dp2)
    report(f"fc_bwd any-batch bn partials B={B}", rel(part.sum(0), ref_part.sum(0)), 1e-5)
    ops.fc_bwd(dl, p2, Wt, None, dp2, lr=0.5)
    report(f"fc_bwd any-batch fused sgd B={B}", rel(Wt, W0.double() - 0.5 * (dl.double().t() @ p2.double())), 1e-6)
