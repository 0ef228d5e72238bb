"""Numerics of the fp32-storage kernel variants (``precision="tf32"``: the reference's own numerics - fp32
tensors, TF32 tensor-core convolutions, fp32 everything else) against plain PyTorch fp32 with TF32 OFF.
Nothing in the oracle is rounded to match the kernels: the tolerance is the TF32 operand rounding itself
(10 explicit mantissa bits) for the tcgen05 convolutions, and fp32 round-off everywhere else."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-12))


def report(name, err, tol):
    print(f"[numerics f32] {name}: rel-L2 {err:.3e} (tol {tol:.1e})")
    assert err < tol, f"{name}: rel-L2 {err:.3e} >= {tol:.1e}"


@pytest.fixture(scope="module")
def ops():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    from torch_distributed_sandbox_b200 import ops as o
    return o


@pytest.fixture(params=[4, 5], ids=["cpasync_planar", "tma_swizzled_rows"])
def feed(ops, request):
    """conv2 forward / dgrad operand feed: 4 = cp.async into 16-byte planes, 5 = one TMA request per row into swizzled rows."""
    ops.conv_tc_set_version(request.param)
    yield request.param
    ops.conv_tc_set_version(5)


@pytest.fixture
def few_ctas(ops):
    ops.set_reserved_sms(ops.sm_count() - 2)
    yield
    ops.set_reserved_sms(0)


def nhwc(x_nchw):
    return x_nchw.permute(0, 2, 3, 1).contiguous()


def nchw(x_nhwc):
    return x_nhwc.permute(0, 3, 1, 2).contiguous()


@pytest.mark.parametrize("shape", [(2, 64, 96), (1, 36, 200), (3, 128, 132), (2, 4, 8)])
def test_conv1_fwd_f32(ops, shape):
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(1)
    x = torch.rand(B, H, W, device=DEV, generator=g)
    w = torch.randn(16, 1, 5, 5, device=DEV, generator=g) * 0.2
    b = torch.randn(16, device=DEV, generator=g) * 0.1
    x1 = torch.full((B, H, W, 16), float("nan"), device=DEV)
    part = torch.zeros(ops.conv1_fwd_grid(), 32, device=DEV)
    n = ops.conv1_fwd_tc(x, w, b, x1, part)
    ref = F.conv2d(x[:, None], w, b, padding=2)
    report(f"conv1_fwd f32 {shape}", rel(nchw(x1), ref), 2e-5)       # hi/lo split: ~16 mantissa bits
    s = part[:n].double().sum(0)
    r = x1.double()
    report("conv1_fwd f32 sum", rel(s[:16], r.sum((0, 1, 2))), 1e-5)
    report("conv1_fwd f32 sumsq", rel(s[16:], (r ** 2).sum((0, 1, 2))), 1e-5)


@pytest.mark.parametrize("C", [16, 32])
def test_bn_relu_pool_f32(ops, C):
    B, H, W = 2, 20, 36
    g = torch.Generator(device=DEV).manual_seed(3)
    x = torch.randn(B, C, H, W, device=DEV, generator=g)
    bn = torch.zeros(4, C, device=DEV)
    bn[0] = torch.randn(C, device=DEV, generator=g)
    bn[1] = torch.randn(C, device=DEV, generator=g) * 0.3
    p = torch.empty(B, H // 2, W // 2, C, device=DEV)
    ops.bn_relu_pool(nhwc(x), bn, p)
    y = x * bn[0].view(1, C, 1, 1) + bn[1].view(1, C, 1, 1)
    report(f"bn_relu_pool f32 C={C}", rel(nchw(p), F.max_pool2d(F.relu(y), 2)), 1e-6)


@pytest.mark.parametrize("shape", [(2, 24, 200), (1, 50, 128), (3, 17, 300), (1, 3, 40), (2, 133, 260)])
def test_conv2_fwd_tf32(ops, shape, feed):
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(4)
    x = torch.randn(B, 16, H, W, device=DEV, generator=g)
    w = torch.randn(32, 16, 5, 5, device=DEV, generator=g) * 0.05
    b = torch.randn(32, device=DEV, generator=g) * 0.1
    out = torch.full((B, H, W, 32), float("nan"), device=DEV)
    part = torch.zeros(ops.conv_tc_max_grid(), 64, device=DEV)
    n = ops.conv2_fwd_tc(nhwc(x), w, b, out, part)
    torch.cuda.synchronize()
    ref = F.conv2d(x, w, b, padding=2)
    report(f"conv2_fwd tf32 {shape}", rel(nchw(out), ref), 1.5e-3)
    r, s = out.double(), part[:n].double().sum(0)
    report("conv2_fwd tf32 sum", rel(s[:32], r.sum((0, 1, 2))), 1e-4)
    report("conv2_fwd tf32 sumsq", rel(s[32:], (r ** 2).sum((0, 1, 2))), 1e-4)


@pytest.mark.parametrize("shape", [(2, 24, 200), (1, 50, 128), (1, 2, 24), (2, 133, 260)])
def test_conv2_dgrad_tf32(ops, shape, feed):
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(5)
    dy = torch.randn(B, 32, H, W, device=DEV, generator=g)
    w = torch.randn(32, 16, 5, 5, device=DEV, generator=g) * 0.05
    out = torch.full((B, H, W, 16), float("nan"), device=DEV)
    ops.conv2_dgrad_tc(nhwc(dy), w, out)
    torch.cuda.synchronize()
    ref = torch.nn.grad.conv2d_input((B, 16, H, W), w, dy, padding=2)
    report(f"conv2_dgrad tf32 {shape}", rel(nchw(out), ref), 1.5e-3)


@pytest.mark.parametrize("shape", [(2, 24, 200), (1, 50, 128), (3, 17, 300), (1, 2, 24), (2, 133, 260)])
def test_conv2_wgrad_tf32(ops, shape, feed):
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(6)
    x = torch.randn(B, 16, H, W, device=DEV, generator=g)
    dy = torch.randn(B, 32, H, W, device=DEV, generator=g)
    part = torch.zeros(ops.conv_tc_max_grid(), 32 * 16 * 25, device=DEV)
    dw = torch.zeros(32, 16, 5, 5, device=DEV)
    ops.conv2_wgrad_tc(nhwc(x), nhwc(dy), part, dw)
    torch.cuda.synchronize()
    ref = torch.nn.grad.conv2d_weight(x, (32, 16, 5, 5), dy, padding=2)
    report(f"conv2_wgrad tf32 {shape}", rel(dw, ref), 1.5e-3)


def test_conv2_tf32_long_items(ops, few_ctas, feed):
    """2 CTAs walk many long items: operand rings and the TMEM ring wrap and carry across items (fp32 slots are twice
    as large, so the ring geometry differs from the bf16 kernels')."""
    B, H, W = 3, 150, 300
    g = torch.Generator(device=DEV).manual_seed(21)
    x = torch.randn(B, 16, H, W, device=DEV, generator=g)
    w = torch.randn(32, 16, 5, 5, device=DEV, generator=g) * 0.05
    b = torch.randn(32, device=DEV, generator=g) * 0.1
    out = torch.full((B, H, W, 32), float("nan"), device=DEV)
    part = torch.zeros(ops.conv_tc_max_grid(), 64, device=DEV)
    n = ops.conv2_fwd_tc(nhwc(x), w, b, out, part)
    torch.cuda.synchronize()
    assert n <= 2
    report("conv2_fwd tf32 long items", rel(nchw(out), F.conv2d(x, w, b, padding=2)), 1.5e-3)
    dy = torch.randn(B, 32, H, W, device=DEV, generator=g)
    dx = torch.full((B, H, W, 16), float("nan"), device=DEV)
    ops.conv2_dgrad_tc(nhwc(dy), w, dx)
    torch.cuda.synchronize()
    report("conv2_dgrad tf32 long items", rel(nchw(dx), torch.nn.grad.conv2d_input((B, 16, H, W), w, dy, padding=2)),
           1.5e-3)
    pw = torch.zeros(ops.conv_tc_max_grid(), 32 * 16 * 25, device=DEV)
    dw = torch.zeros(32, 16, 5, 5, device=DEV)
    ops.conv2_wgrad_tc(nhwc(x), nhwc(dy), pw, dw)
    torch.cuda.synchronize()
    report("conv2_wgrad tf32 long items", rel(dw, torch.nn.grad.conv2d_weight(x, (32, 16, 5, 5), dy, padding=2)),
           1.5e-3)


@pytest.mark.parametrize("B", [1, 5, 8])
def test_fc_f32(ops, B):
    Fdim = 32 * 12 * 20
    g = torch.Generator(device=DEV).manual_seed(7)
    p2 = torch.rand(B, Fdim, device=DEV, generator=g)
    W = torch.randn(10, Fdim, device=DEV, generator=g) * 0.02
    bias = torch.randn(10, device=DEV, generator=g) * 0.1
    part = torch.zeros(ops.fc_grid(), 80, device=DEV)
    logits = torch.zeros(B, 10, device=DEV)
    ops.fc_fwd(p2, W, bias, part, logits)
    report(f"fc_fwd f32 B={B}", rel(logits, p2.double() @ W.double().t() + bias.double()), 1e-6)
    dl = torch.randn(B, 10, device=DEV, generator=g) * 0.1
    dW, dp2, W0 = torch.zeros_like(W), torch.empty_like(p2), W.clone()
    ops.fc_bwd(dl, p2, W, dW, dp2)
    report("fc_bwd f32 dW", rel(dW, dl.double().t() @ p2.double()), 1e-6)
    report("fc_bwd f32 dp2", rel(dp2, dl.double() @ W0.double()), 1e-6)
    ops.fc_bwd(dl, p2, W, None, dp2, lr=0.5)
    report("fc_bwd f32 fused sgd", rel(W, W0.double() - 0.5 * (dl.double().t() @ p2.double())), 1e-6)


@pytest.mark.parametrize("C", [16, 32])
def test_pool_bn_bwd_f32(ops, C):
    B, H, W = 2, 24, 40
    g = torch.Generator(device=DEV).manual_seed(8)
    x = (torch.randn(B, C, H, W, device=DEV, generator=g) * 1.5 + 0.3).requires_grad_(True)
    dp = torch.randn(B, C, H // 2, W // 2, device=DEV, generator=g)
    gamma = torch.rand(C, device=DEV, generator=g) + 0.5
    gamma[::3] *= -1
    beta = torch.randn(C, device=DEV, generator=g) * 0.2
    gam, bet = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    y = F.batch_norm(x, None, None, gam, bet, training=True, eps=1e-5)
    F.max_pool2d(F.relu(y), 2).backward(dp)
    xd = x.detach()
    mean, var = xd.mean((0, 2, 3)), xd.var((0, 2, 3), unbiased=False)
    rstd = 1 / torch.sqrt(var + 1e-5)
    bn = torch.stack([gamma * rstd, beta - mean * gamma * rstd, mean, rstd]).contiguous()
    grid = ops.pool_bn_bwd_grid()
    part, red = torch.zeros(grid, 64, device=DEV), torch.zeros(2, C, device=DEV)
    dg, db = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    xh, dph = nhwc(xd), nhwc(dp)
    ops.pool_bn_bwd(xh, dph, bn, red, part, None, mode=0)
    ops.bn_bwd_finalize(part, grid, C, B * H * W, red, dg, db)
    report(f"bn bwd f32 dgamma C={C}", rel(dg, gam.grad), 1e-4)
    report(f"bn bwd f32 dbeta C={C}", rel(db, bet.grad), 1e-4)
    dx = torch.empty_like(xh)
    ops.pool_bn_bwd(xh, dph, bn, red, None, dx, mode=1)
    report(f"pool_bn_bwd f32 dx C={C}", rel(nchw(dx), x.grad), 1e-4)


@pytest.mark.parametrize("impl", ["tc", "ffma"])
@pytest.mark.parametrize("shape", [(2, 40, 72), (1, 8, 128), (3, 36, 300), (2, 132, 260)])
def test_conv1_wgrad_f32(ops, shape, impl):
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(9)
    xin = torch.rand(B, H, W, device=DEV, generator=g)
    w = torch.randn(16, 1, 5, 5, device=DEV, generator=g) * 0.2
    b = torch.zeros(16, device=DEV)
    gamma = torch.rand(16, device=DEV, generator=g) + 0.5
    beta = torch.randn(16, device=DEV, generator=g) * 0.2
    dp = nhwc(torch.randn(B, 16, H // 2, W // 2, device=DEV, generator=g))
    x1 = torch.empty(B, H, W, 16, device=DEV)
    part = torch.zeros(ops.conv1_fwd_grid(), 32, device=DEV)
    n = ops.conv1_fwd_tc(xin, w, b, x1, part)
    bn = torch.zeros(4, 16, device=DEV)
    rm, rv = torch.zeros(16, device=DEV), torch.ones(16, device=DEV)
    ops.bn_finalize(part, n, 16, B * H * W, gamma, beta, rm, rv, None, bn)
    grid = ops.pool_bn_bwd_grid()
    partb, red = torch.zeros(grid, 64, device=DEV), torch.zeros(2, 16, device=DEV)
    dg, db = torch.zeros(16, device=DEV), torch.zeros(16, device=DEV)
    ops.pool_bn_bwd(x1, dp, bn, red, partb, None, mode=0)
    ops.bn_bwd_finalize(partb, grid, 16, B * H * W, red, dg, db)
    pw = torch.zeros(ops.conv1_wgrad_grid(), 400, device=DEV)
    dw = torch.zeros(16, 1, 5, 5, device=DEV)
    (ops.conv1_wgrad_tc if impl == "tc" else ops.conv1_wgrad)(xin, x1, dp, bn, red, pw, dw)
    x1r = nchw(x1).requires_grad_(True)           # oracle on the same stored conv output (arg-max ties)
    y = F.batch_norm(x1r, None, None, gamma, beta, training=True)
    F.max_pool2d(F.relu(y), 2).backward(nchw(dp))
    ref = torch.nn.grad.conv2d_weight(xin[:, None], (16, 1, 5, 5), x1r.grad, padding=2)
    # tc: g and the input are rounded to TF32 (10-bit mantissa), fp32 accumulate; ffma: exact fp32 products
    report(f"conv1_wgrad f32[{impl}] {shape}", rel(dw, ref), 2e-3)


@pytest.mark.parametrize("shape", [(2, 24, 200), (1, 50, 128)])
def test_conv2_dgrad_fused_bn1_stats_tf32(ops, shape, feed):
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(22)
    x1 = nhwc(torch.randn(B, 16, 2 * H, 2 * W, device=DEV, generator=g) * 1.3 + 0.2)
    gamma = torch.rand(16, device=DEV, generator=g) + 0.5
    gamma[1::4] *= -1
    beta = torch.randn(16, device=DEV, generator=g) * 0.3
    mean, var = x1.mean((0, 1, 2)), x1.var((0, 1, 2), unbiased=False)
    rstd = 1 / torch.sqrt(var + 1e-5)
    bn1 = torch.stack([gamma * rstd, beta - mean * gamma * rstd, mean, rstd]).contiguous()
    p1 = torch.empty(B, H, W, 16, device=DEV)
    ops.bn_relu_pool(x1, bn1, p1)
    dy = nhwc(torch.randn(B, 32, H, W, device=DEV, generator=g))
    w = torch.randn(32, 16, 5, 5, device=DEV, generator=g) * 0.05
    ref_dp1 = torch.empty(B, H, W, 16, device=DEV)
    ops.conv2_dgrad_tc(dy, w, ref_dp1)
    part = torch.zeros(ops.conv_tc_max_grid(), 32, device=DEV)
    inplace = p1.clone()
    n = ops.conv2_dgrad_stats_tc(dy, w, inplace, bn1, inplace, part)
    assert torch.equal(inplace, ref_dp1)
    red_f, dg_f, db_f = torch.zeros(2, 16, device=DEV), torch.zeros(16, device=DEV), torch.zeros(16, device=DEV)
    ops.bn_bwd_finalize(part, n, 16, B * 4 * H * W, red_f, dg_f, db_f)
    grid = ops.pool_bn_bwd_grid()
    part2 = torch.zeros(grid, 64, device=DEV)
    red_r, dg_r, db_r = torch.zeros(2, 16, device=DEV), torch.zeros(16, device=DEV), torch.zeros(16, device=DEV)
    ops.pool_bn_bwd(x1, ref_dp1, bn1, red_r, part2, None, mode=0)
    ops.bn_bwd_finalize(part2, grid, 16, B * 4 * H * W, red_r, dg_r, db_r)
    report(f"fused bn1 bwd f32 dbeta {shape}", rel(db_f, db_r), 1e-5)
    report(f"fused bn1 bwd f32 dgamma {shape}", rel(dg_f, dg_r), 1e-3)


@pytest.mark.parametrize("shape", [(2, 64, 96), (1, 36, 200), (3, 128, 132), (2, 4, 8)])
def test_conv1_two_pass_equals_conv_then_pool(ops, shape):
    """Statistics-only pass + conv/BN/ReLU/pool pass == conv1_fwd_tc followed by bn_relu_pool, bit for bit."""
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(41)
    x = torch.rand(B, H, W, device=DEV, generator=g)
    w = torch.randn(16, 1, 5, 5, device=DEV, generator=g) * 0.2
    b = torch.randn(16, device=DEV, generator=g) * 0.1
    bn = torch.zeros(4, 16, device=DEV)
    bn[0] = torch.randn(16, device=DEV, generator=g)          # scale, both signs
    bn[1] = torch.randn(16, device=DEV, generator=g) * 0.3
    x1a = torch.empty(B, H, W, 16, device=DEV)
    pa = torch.zeros(ops.conv1_fwd_grid(), 32, device=DEV)
    na = ops.conv1_fwd_tc(x, w, b, x1a, pa)
    p1a = torch.empty(B, H // 2, W // 2, 16, device=DEV)
    ops.bn_relu_pool(x1a, bn, p1a)
    pb = torch.zeros(ops.conv1_fwd_grid(), 32, device=DEV)
    nb = ops.conv1_stats_tc(x, w, b, pb)
    assert na == nb and torch.equal(pa[:na], pb[:nb])
    x1b = torch.full((B, H, W, 16), float("nan"), device=DEV)
    p1b = torch.full((B, H // 2, W // 2, 16), float("nan"), device=DEV)
    ops.conv1_fwd_pool_tc(x, w, b, bn, x1b, p1b, torch.zeros(ops.conv1_fwd_grid(), 32, device=DEV))
    torch.cuda.synchronize()
    assert torch.equal(x1a, x1b)
    assert torch.equal(p1a, p1b), float((p1a - p1b).abs().max())


@pytest.mark.parametrize("shape", [(2, 40, 72), (1, 8, 128), (3, 64, 300), (2, 136, 260)])
def test_conv1_wgrad_without_the_conv_output(ops, shape):
    """Forward pass 2 emits arg-max/ReLU bytes instead of (or next to) x1; the x1-free weight gradient (routed part on
    the tensor cores + dense BatchNorm term from the input's Gram matrix) must match the fp32 PyTorch oracle."""
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(43)
    xin = torch.rand(B, H, W, device=DEV, generator=g)
    w = torch.randn(16, 1, 5, 5, device=DEV, generator=g) * 0.2
    b = torch.randn(16, device=DEV, generator=g) * 0.05
    gamma = torch.rand(16, device=DEV, generator=g) + 0.5
    gamma[::5] *= -1
    beta = torch.randn(16, device=DEV, generator=g) * 0.2
    dp = nhwc(torch.randn(B, 16, H // 2, W // 2, device=DEV, generator=g))
    part = torch.zeros(ops.conv1_fwd_grid(), 32, device=DEV)
    n = ops.conv1_stats_tc(xin, w, b, part)
    bn = torch.zeros(4, 16, device=DEV)
    rm, rv = torch.zeros(16, device=DEV), torch.ones(16, device=DEV)
    ops.bn_finalize(part, n, 16, B * H * W, gamma, beta, rm, rv, None, bn)
    x1 = torch.empty(B, H, W, 16, device=DEV)
    p1 = torch.empty(B, H // 2, W // 2, 16, device=DEV)
    mask = torch.full((B, H // 2, W // 2, 16), 255, device=DEV, dtype=torch.uint8)
    ops.conv1_fwd_pool_tc(xin, w, b, bn, x1, p1, torch.zeros_like(part), amask_u8=mask)
    p1_ref = torch.empty_like(p1)
    ops.bn_relu_pool(x1, bn, p1_ref)
    assert torch.equal(p1, p1_ref)
    # the mask agrees with torch's max-pool indices wherever the window maximum is unique and positive
    y = nchw(x1) * bn[0].view(1, 16, 1, 1) + bn[1].view(1, 16, 1, 1)
    pooled, idx = F.max_pool2d(y, 2, return_indices=True)
    yy, xx = idx // W, idx % W
    am_ref = ((yy % 2) * 2 + (xx % 2)).permute(0, 2, 3, 1)
    relu_ref = (pooled > 0).permute(0, 2, 3, 1)
    assert torch.equal((mask & 3).long(), am_ref) and torch.equal((mask >> 2).bool(), relu_ref)
    # BN backward reductions (from x1, as the engine gets them from the dgrad epilogue), then both weight gradients
    grid = ops.pool_bn_bwd_grid()
    partb, red = torch.zeros(grid, 64, device=DEV), torch.zeros(2, 16, device=DEV)
    dg, db = torch.zeros(16, device=DEV), torch.zeros(16, device=DEV)
    ops.pool_bn_bwd(x1, dp, bn, red, partb, None, mode=0)
    ops.bn_bwd_finalize(partb, grid, 16, B * H * W, red, dg, db)
    pw = torch.zeros(ops.conv1_wgrad_grid(), ops.conv1_wgrad_nox1_partial(), device=DEV)
    dw = torch.zeros(16, 1, 5, 5, device=DEV)
    ops.conv1_wgrad_nox1(xin, dp, mask, w, b, bn, red, pw, dw)
    x1r = nchw(x1).requires_grad_(True)
    yb = F.batch_norm(x1r, None, None, gamma, beta, training=True)
    F.max_pool2d(F.relu(yb), 2).backward(nchw(dp))
    ref = torch.nn.grad.conv2d_weight(xin[:, None], (16, 1, 5, 5), x1r.grad, padding=2)
    report(f"conv1_wgrad without x1 {shape}", rel(dw, ref), 3e-3)


@pytest.mark.parametrize("shape", [(2, 40, 72), (1, 8, 128), (3, 64, 300), (2, 136, 1100), (1, 302, 2020)])
def test_conv1_row_operand_two_pass(ops, shape):
    """csrc/conv1_hz.cu: the kind::tf32 layer-1 forward whose MMA operand is the image row as TMA delivered it.
    Pass 1 = batch statistics, pass 2 = pooled tensor + arg-max/ReLU bytes; both against plain fp32 PyTorch
    (TF32 error level: the input is truncated to 10 mantissa bits, as cuDNN's TF32 conv of the reference does)."""
    B, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(47)
    xin = torch.rand(B, H, W, device=DEV, generator=g)
    w = torch.randn(16, 1, 5, 5, device=DEV, generator=g) * 0.2
    b = torch.randn(16, device=DEV, generator=g) * 0.1
    gamma = torch.rand(16, device=DEV, generator=g) + 0.5
    gamma[::5] *= -1
    beta = torch.randn(16, device=DEV, generator=g) * 0.2
    x1 = F.conv2d(xin[:, None], w, b, padding=2)
    part = torch.zeros(ops.conv1_fwd_grid(), 32, device=DEV)
    n = ops.conv1_stats_tc(xin, w, b, part, row_operand=True)
    sums = part[:n].double().sum(0)
    ref_sum, ref_sq = x1.double().sum((0, 2, 3)), (x1.double() ** 2).sum((0, 2, 3))
    report(f"conv1 row-operand stats: sum {shape}", rel(sums[:16].float(), ref_sum.float()), 2e-3)
    report(f"conv1 row-operand stats: sumsq {shape}", rel(sums[16:].float(), ref_sq.float()), 2e-3)
    bn = torch.zeros(4, 16, device=DEV)
    rm, rv = torch.zeros(16, device=DEV), torch.ones(16, device=DEV)
    ops.bn_finalize(part, n, 16, B * H * W, gamma, beta, rm, rv, None, bn)
    p1 = torch.full((B, H // 2, W // 2, 16), float("nan"), device=DEV)
    mask = torch.full((B, H // 2, W // 2, 16), 255, device=DEV, dtype=torch.uint8)
    ops.conv1_fwd_pool_tc(xin, w, b, bn, None, p1, None, amask_u8=mask, row_operand=True)
    torch.cuda.synchronize()
    y = x1 * bn[0].view(1, 16, 1, 1) + bn[1].view(1, 16, 1, 1)
    pooled, idx = F.max_pool2d(y, 2, return_indices=True)
    report(f"conv1 row-operand pooled {shape}", rel(nchw(p1), F.relu(pooled)), 2e-3)
    # arg-max / ReLU bytes: equal to the fp32 oracle's except where TF32 rounding flips a near-tie
    yy, xx = idx // W, idx % W
    am_ref = ((yy % 2) * 2 + (xx % 2)).permute(0, 2, 3, 1)
    relu_ref = (pooled > 0).permute(0, 2, 3, 1)
    assert int(mask.max()) <= 7
    bad_am = ((mask & 3).long() != am_ref).float().mean().item()
    bad_relu = ((mask >> 2).bool() != relu_ref).float().mean().item()
    print(f"[f32] row-operand mask mismatch vs fp32 oracle: arg-max {bad_am:.2e}, relu {bad_relu:.2e}")
    assert bad_am < 5e-3 and bad_relu < 5e-3
    # Against an oracle that rounds its operands the way the kernel does (input truncated, weights rounded to nearest,
    # exact accumulation) the choice of arg-max is the same up to accumulation-order ties, and the weight gradient
    # routed with the kernel's bytes matches to TF32 error.  (Against the plain fp32 oracle a flipped near-tie moves a
    # gradient value by one pixel - on this white-noise input that alone is a few percent of dW.)
    xt = (xin.view(torch.int32) & -8192).view(torch.float32)
    wt = ((w.view(torch.int32) + 4096) & -8192).view(torch.float32)
    x1t = (F.conv2d(xt[:, None].double(), wt.double(), b.double(), padding=2)).float()
    yt = x1t * bn[0].view(1, 16, 1, 1) + bn[1].view(1, 16, 1, 1)
    pooled_t, idx_t = F.max_pool2d(yt, 2, return_indices=True)
    am_t = (((idx_t // W) % 2) * 2 + (idx_t % W) % 2).permute(0, 2, 3, 1)
    bad_t = ((mask & 3).long() != am_t).float().mean().item()
    bad_rt = ((mask >> 2).bool() != (pooled_t > 0).permute(0, 2, 3, 1)).float().mean().item()
    print(f"[f32] row-operand mask mismatch vs operand-rounded oracle: arg-max {bad_t:.2e}, relu {bad_rt:.2e}")
    assert bad_t < 2e-5 and bad_rt < 2e-5
    dp = nhwc(torch.randn(B, 16, H // 2, W // 2, device=DEV, generator=g))
    if H % 8 == 0:
        x1r = x1t.clone().requires_grad_(True)
        yb = F.batch_norm(x1r, None, None, gamma, beta, training=True)
        F.max_pool2d(F.relu(yb), 2).backward(nchw(dp))
        ref = torch.nn.grad.conv2d_weight(xin[:, None], (16, 1, 5, 5), x1r.grad, padding=2)
        yd = yb.detach().requires_grad_(True)
        F.max_pool2d(F.relu(yd), 2).backward(nchw(dp))
        xhat = (x1t - bn[2].view(1, 16, 1, 1)) * bn[3].view(1, 16, 1, 1)
        red = torch.stack([yd.grad.mean((0, 2, 3)), (yd.grad * xhat).mean((0, 2, 3))]).contiguous()
        pw = torch.zeros(ops.conv1_wgrad_grid(), ops.conv1_wgrad_nox1_partial(), device=DEV)
        dw = torch.zeros(16, 1, 5, 5, device=DEV)
        ops.conv1_wgrad_nox1(xin, dp, mask, w, b, bn, red, pw, dw)
        report(f"conv1_wgrad from row-operand mask {shape}", rel(dw, ref), 5e-3)
