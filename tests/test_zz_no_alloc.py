"""The allocation-free conv1 weight-gradient entry point, driven through the C ABI.

Who needs it: a caller that may not allocate device memory at the point where it wants this
gradient - it is inside a CUDA stream capture or a "no new allocations" memory-pool phase, it was
given a fixed device-memory budget that is already spent, or it is a thin C/Fortran driver
that owns nothing but its input and output arrays.  ``b200_conv1_wgrad`` asks such a caller for
[grid][400] floats of scratch; ``b200_conv1_wgrad_noalloc`` keeps that scratch in a 256,000-byte
array compiled into the library instead.

Checked here, calling both entry points with ctypes exactly as a C program would:
  * the result equals the caller-scratch entry point's (same grid: the same bits; a different
    grid: within the band the other kernel tests use), on two calls in a row with different
    inputs - the second call sees the first call's leftovers in the shared scratch;
  * it also equals a plain PyTorch fp32 computation of the same gradient;
  * a grid larger than the built-in scratch is refused with a message, nothing is launched.
The file name sorts last on purpose: a run that stops at the first failure has then already
run every other test.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes
from ctypes import c_float, c_int, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _lib():
    from torch_distributed_sandbox_b200 import _native
    lib = _native.cuda()
    lib.b200_last_error.restype = ctypes.c_char_p
    lib.b200_conv1_wgrad.argtypes = [c_void_p] * 7 + [c_int, c_int, c_int, c_float, c_void_p]
    lib.b200_conv1_wgrad_noalloc.argtypes = [c_void_p] * 6 + [c_int, c_int, c_int, c_float, c_int, c_void_p]
    return lib


def _inputs(seed, B, H, W):
    g = torch.Generator(device=DEV).manual_seed(seed)
    xin = torch.rand(B, H, W, device=DEV, generator=g)
    x1 = (torch.randn(B, H, W, 16, device=DEV, generator=g) * 0.7).to(torch.bfloat16)
    dp1 = torch.randn(B, H // 2, W // 2, 16, device=DEV, generator=g).to(torch.bfloat16)
    bn = torch.empty(4, 16, device=DEV)
    bn[0] = torch.rand(16, device=DEV, generator=g) + 0.5          # scale
    bn[1] = torch.randn(16, device=DEV, generator=g) * 0.2          # shift
    bn[2] = torch.randn(16, device=DEV, generator=g) * 0.1          # mean
    bn[3] = torch.rand(16, device=DEV, generator=g) + 0.8           # rstd
    red = torch.randn(2, 16, device=DEV, generator=g) * 0.01
    return xin, x1, dp1, bn, red


def rel(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-12))


def _torch_oracle(xin, x1, dp1, bn, red):
    """Plain PyTorch fp32 statement of what the kernel computes: the pool / ReLU / BatchNorm backward of dp1 through
    the stored conv output x1 (first maximum of each 2x2 window, as ATen picks it), then the conv1 weight gradient."""
    import torch.nn.functional as F
    x = x1.float().permute(0, 3, 1, 2)                      # [B,16,H,W]
    g = dp1.float().permute(0, 3, 1, 2)
    sc, sh, mu, rs = (bn[i].view(1, 16, 1, 1) for i in range(4))
    mdy, mdyx = red[0].view(1, 16, 1, 1), red[1].view(1, 16, 1, 1)
    y = x * sc + sh
    B, C, H, W = y.shape
    win = y.view(B, C, H // 2, 2, W // 2, 2).permute(0, 1, 2, 4, 3, 5).reshape(B, C, H // 2, W // 2, 4)
    best, am = win.max(dim=-1)                              # torch.max returns the first maximum on ties
    routed = torch.where(best > 0, g, torch.zeros_like(g))
    onehot = F.one_hot(am, 4).to(routed.dtype) * routed.unsqueeze(-1)
    dense = onehot.view(B, C, H // 2, W // 2, 2, 2).permute(0, 1, 2, 4, 3, 5).reshape(B, C, H, W)
    dx = sc * (dense - mdy - (x - mu) * rs * mdyx)
    return torch.nn.grad.conv2d_weight(xin[:, None], (16, 1, 5, 5), dx, padding=2)


def test_noalloc_matches_caller_scratch_entry_point_twice_and_rejects_oversize():
    lib = _lib()
    stream = c_void_p(torch.cuda.current_stream().cuda_stream)
    nsm = int(lib.b200_conv1_wgrad_grid())
    cap = int(lib.b200_conv1_wgrad_noalloc_max_blocks())
    assert cap >= nsm, "the built-in scratch must cover one block per SM on this GPU"
    for call, (seed, shape) in enumerate([(11, (2, 64, 96)), (12, (3, 40, 72))]):
        B, H, W = shape
        xin, x1, dp1, bn, red = _inputs(seed, B, H, W)
        ptr = lambda t: c_void_p(t.data_ptr())
        # existing entry point: caller-provided scratch
        scratch = torch.empty(nsm, 400, device=DEV)
        want = torch.zeros(16, 1, 5, 5, device=DEV)
        rc = lib.b200_conv1_wgrad(ptr(xin), ptr(x1), ptr(dp1), ptr(bn), ptr(red), ptr(scratch), ptr(want), B, H, W,
                                  1.0, stream)
        assert rc == 0, lib.b200_last_error()
        # new entry point, same grid (nblocks = 0 -> SM count): the same arithmetic in the same order
        got = torch.full((16, 1, 5, 5), float("nan"), device=DEV)
        rc = lib.b200_conv1_wgrad_noalloc(ptr(xin), ptr(x1), ptr(dp1), ptr(bn), ptr(red), ptr(got), B, H, W, 1.0, 0,
                                          stream)
        assert rc == 0, lib.b200_last_error()
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
        assert torch.equal(got, want), f"call {call}: rel-L2 {rel(got, want):.3e}"
        # ... and both agree with an independent fp32 PyTorch computation of the same quantity
        oracle = _torch_oracle(xin, x1, dp1, bn, red)
        assert rel(got, oracle) < 2e-3, f"call {call}: rel-L2 vs torch {rel(got, oracle):.3e}"
        # a different grid re-partitions the sum: equal within the kernel tests' tolerance
        got2 = torch.zeros(16, 1, 5, 5, device=DEV)
        rc = lib.b200_conv1_wgrad_noalloc(ptr(xin), ptr(x1), ptr(dp1), ptr(bn), ptr(red), ptr(got2), B, H, W, 1.0,
                                          37, stream)
        assert rc == 0, lib.b200_last_error()
        torch.cuda.synchronize()
        assert rel(got2, want) < 2e-3, f"call {call}: rel-L2 {rel(got2, want):.3e}"
    # beyond the built-in scratch: refused, with a message, and the output is left untouched
    untouched = torch.full((16, 1, 5, 5), 7.0, device=DEV)
    rc = lib.b200_conv1_wgrad_noalloc(ptr(xin), ptr(x1), ptr(dp1), ptr(bn), ptr(red), ptr(untouched), B, H, W, 1.0,
                                      cap + 1, stream)
    assert rc != 0
    msg = lib.b200_last_error().decode()
    assert "built-in scratch" in msg and str(cap) in msg, msg
    torch.cuda.synchronize()
    assert bool((untouched == 7.0).all())
