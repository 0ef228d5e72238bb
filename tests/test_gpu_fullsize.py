"""One full-size (3000x3000 input, 1500x1500 at conv2) single-image check per tensor-core kernel against plain
PyTorch fp32 with TF32 off.  At this size every persistent CTA walks many long work
items - the production geometry: operand rings and TMEM accumulator rings wrap hundreds of times, which the small
shapes of the other kernel tests only reach with the `few_ctas` fixture.  The oracle's operands are NOT rounded to
match: the bands are the storage / operand precision of each mode (bf16: 8 mantissa bits, tf32: 11)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
S = 3000


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-12))


def report(name, err, tol):
    print(f"[fullsize] {name}: rel-L2 {err:.3e} (tol {tol:.1e})")
    assert err < tol, f"{name}: rel-L2 {err:.3e} >= {tol:.1e}"


@pytest.fixture(scope="module")
def ops():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    from torch_distributed_sandbox_b200 import ops as o
    return o


def nhwc(x, dt):
    return x.permute(0, 2, 3, 1).contiguous().to(dt)


def nchw(x):
    return x.float().permute(0, 3, 1, 2).contiguous()


MODES = [(torch.bfloat16, 1e-2, "bf16"), (torch.float32, 2e-3, "tf32")]


@pytest.mark.parametrize("dt,tol,name", MODES)
def test_conv1_fwd_fullsize(ops, dt, tol, name):
    g = torch.Generator(device=DEV).manual_seed(31)
    x = torch.rand(1, S, S, device=DEV, generator=g)
    w = torch.randn(16, 1, 5, 5, device=DEV, generator=g) * 0.2
    b = torch.randn(16, device=DEV, generator=g) * 0.1
    x1 = torch.full((1, S, S, 16), float("nan"), device=DEV, dtype=dt)
    part = torch.zeros(ops.conv1_fwd_grid(), 32, device=DEV)
    n = ops.conv1_fwd_tc(x, w, b, x1, part)
    torch.cuda.synchronize()
    report(f"conv1_fwd {name}", rel(nchw(x1), F.conv2d(x[:, None], w, b, padding=2)), 4e-3 if dt == torch.bfloat16 else 2e-5)
    s = part[:n].double().sum(0)
    report(f"conv1_fwd {name} sum", rel(s[:16], x1.double().sum((0, 1, 2))), 1e-5)


@pytest.mark.parametrize("dt,tol,name", MODES)
def test_conv2_fwd_dgrad_wgrad_fullsize(ops, dt, tol, name):
    H = W = S // 2
    g = torch.Generator(device=DEV).manual_seed(32)
    x = torch.randn(1, 16, H, W, device=DEV, generator=g)
    w = torch.randn(32, 16, 5, 5, device=DEV, generator=g) * 0.05
    b = torch.randn(32, device=DEV, generator=g) * 0.1
    xs = nhwc(x, dt)
    out = torch.full((1, H, W, 32), float("nan"), device=DEV, dtype=dt)
    part = torch.zeros(ops.conv_tc_max_grid(), 64, device=DEV)
    n = ops.conv2_fwd_tc(xs, w, b, out, part)
    torch.cuda.synchronize()
    report(f"conv2_fwd {name}", rel(nchw(out), F.conv2d(x, w, b, padding=2)), tol)
    s = part[:n].double().sum(0)
    report(f"conv2_fwd {name} sum", rel(s[:32], out.double().sum((0, 1, 2))), 1e-4)
    dy = torch.randn(1, 32, H, W, device=DEV, generator=g)
    dys = nhwc(dy, dt)
    dx = torch.full((1, H, W, 16), float("nan"), device=DEV, dtype=dt)
    ops.conv2_dgrad_tc(dys, w, dx)
    torch.cuda.synchronize()
    report(f"conv2_dgrad {name}", rel(nchw(dx), torch.nn.grad.conv2d_input((1, 16, H, W), w, dy, padding=2)), tol)
    pw = torch.zeros(ops.conv_tc_max_grid(), 32 * 16 * 25, device=DEV)
    dw = torch.zeros(32, 16, 5, 5, device=DEV)
    ops.conv2_wgrad_tc(xs, dys, pw, dw)
    torch.cuda.synchronize()
    report(f"conv2_wgrad {name}", rel(dw, torch.nn.grad.conv2d_weight(x, (32, 16, 5, 5), dy, padding=2)), tol)


@pytest.mark.parametrize("dt,tol,name", MODES)
def test_conv1_wgrad_fullsize(ops, dt, tol, name):
    g = torch.Generator(device=DEV).manual_seed(33)
    xin = torch.rand(1, S, S, device=DEV, generator=g)
    w = torch.randn(16, 1, 5, 5, device=DEV, generator=g) * 0.2
    gamma = torch.rand(16, device=DEV, generator=g) + 0.5
    beta = torch.randn(16, device=DEV, generator=g) * 0.2
    dp = nhwc(torch.randn(1, 16, S // 2, S // 2, device=DEV, generator=g), dt)
    x1 = torch.empty(1, S, S, 16, device=DEV, dtype=dt)
    part = torch.zeros(ops.conv1_fwd_grid(), 32, device=DEV)
    n = ops.conv1_fwd_tc(xin, w, torch.zeros(16, device=DEV), x1, part)
    bn = torch.zeros(4, 16, device=DEV)
    rm, rv = torch.zeros(16, device=DEV), torch.ones(16, device=DEV)
    ops.bn_finalize(part, n, 16, S * S, gamma, beta, rm, rv, None, bn)
    grid = ops.pool_bn_bwd_grid()
    partb, red = torch.zeros(grid, 64, device=DEV), torch.zeros(2, 16, device=DEV)
    dg, db = torch.zeros(16, device=DEV), torch.zeros(16, device=DEV)
    ops.pool_bn_bwd(x1, dp, bn, red, partb, None, mode=0)
    ops.bn_bwd_finalize(partb, grid, 16, S * S, red, dg, db)
    pw = torch.zeros(ops.conv1_wgrad_grid(), 400, device=DEV)
    dw = torch.zeros(16, 1, 5, 5, device=DEV)
    ops.conv1_wgrad_tc(xin, x1, dp, bn, red, pw, dw)
    torch.cuda.synchronize()
    # oracle on the stored conv output (identical pooling arg-max), fp32 everywhere else
    x1r = nchw(x1).requires_grad_(True)
    y = F.batch_norm(x1r, None, None, gamma, beta, training=True)
    F.max_pool2d(F.relu(y), 2).backward(nchw(dp))
    ref = torch.nn.grad.conv2d_weight(xin[:, None], (16, 1, 5, 5), x1r.grad, padding=2)
    # 9 M pixels of BatchNorm-backward cancellation amplify the operand rounding of each mode: the bf16 kernel
    # (bf16 gradient and input operands) is measured at 8e-2 here, the tf32 one at ~1e-3
    report(f"conv1_wgrad {name}", rel(dw, ref), 1.5e-1 if dt == torch.bfloat16 else 5e-3)


@pytest.mark.parametrize("row_operand", [True, False])
def test_conv1_two_pass_and_wgrad_without_x1_fullsize(ops, row_operand):
    """The tf32 mode's layer 1 at 3000x3000: statistics pass, conv+BN+ReLU+pool pass that stores no conv output (only the
    pooled tensor and the arg-max/ReLU bytes), weight gradient from those bytes + the input's Gram matrix - against
    plain fp32 PyTorch on the explicitly materialised conv output."""
    g = torch.Generator(device=DEV).manual_seed(34)
    xin = torch.rand(1, S, S, device=DEV, generator=g)
    w = torch.randn(16, 1, 5, 5, device=DEV, generator=g) * 0.2
    b = torch.randn(16, device=DEV, generator=g) * 0.05
    gamma = torch.rand(16, device=DEV, generator=g) + 0.5
    beta = torch.randn(16, device=DEV, generator=g) * 0.2
    dp = nhwc(torch.randn(1, 16, S // 2, S // 2, device=DEV, generator=g), torch.float32)
    part = torch.zeros(ops.conv1_fwd_grid(), 32, device=DEV)
    n = ops.conv1_stats_tc(xin, w, b, part, row_operand=row_operand)
    bn = torch.zeros(4, 16, device=DEV)
    rm, rv = torch.zeros(16, device=DEV), torch.ones(16, device=DEV)
    ops.bn_finalize(part, n, 16, S * S, gamma, beta, rm, rv, None, bn)
    p1 = torch.empty(1, S // 2, S // 2, 16, device=DEV)
    mask = torch.empty(1, S // 2, S // 2, 16, device=DEV, dtype=torch.uint8)
    ops.conv1_fwd_pool_tc(xin, w, b, bn, None, p1, torch.zeros_like(part), amask_u8=mask, row_operand=row_operand)
    if row_operand:      # kind::tf32 kernel: the oracle rounds its operands the same way (input truncated, weights rna)
        xo = (xin.view(torch.int32) & -8192).view(torch.float32)
        wo = ((w.view(torch.int32) + 4096) & -8192).view(torch.float32)
    else:
        xo, wo = xin, w
    x1 = F.conv2d(xo[:, None], wo, b, padding=2)                        # oracle conv output, fp32
    x1r = x1.clone().requires_grad_(True)
    y = F.batch_norm(x1r, None, None, gamma, beta, training=True)
    pooled = F.max_pool2d(F.relu(y), 2)
    report("conv1 two-pass pooled output", rel(nchw(p1), pooled.detach()), 2e-5 if not row_operand else 1e-4)
    pooled.backward(nchw(dp))
    ref = torch.nn.grad.conv2d_weight(xin[:, None], (16, 1, 5, 5), x1r.grad, padding=2)
    # BN backward reductions from the oracle (the engine takes them from the dgrad epilogue)
    xhat = (x1 - bn[2].view(1, 16, 1, 1)) * bn[3].view(1, 16, 1, 1)
    yd = y.detach().requires_grad_(True)
    F.max_pool2d(F.relu(yd), 2).backward(nchw(dp))
    dy = yd.grad
    red = torch.stack([dy.mean((0, 2, 3)), (dy * xhat).mean((0, 2, 3))]).contiguous()
    pw = torch.zeros(ops.conv1_wgrad_grid(), ops.conv1_wgrad_nox1_partial(), device=DEV)
    dw = torch.zeros(16, 1, 5, 5, device=DEV)
    ops.conv1_wgrad_nox1(xin, dp, mask, w, b, bn, red, pw, dw)
    torch.cuda.synchronize()
    report("conv1_wgrad without x1, full size", rel(dw, ref), 5e-3)
