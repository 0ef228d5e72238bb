"""Python front-ends of the hand-written sm_100a kernels (C ABI in ``include/b200dp.h``).

Every function takes torch CUDA tensors, launches on torch's *current* stream and
returns nothing (outputs are pre-allocated by the caller: the engine owns all
activation memory so a training step allocates nothing).  ``LAUNCHES`` counts
kernels launched by this module; ``bench.py`` reports it as ``gpu_launches``.

Layouts: activations are **NHWC**, bf16 (fast mode) or fp32 (the reference's precision:
fp32 tensors, TF32 tensor-core math in the convolutions); the storage dtype of the
activation argument selects the kernel variant (``*_f32`` entry points).  Weights/grads
are fp32 in the reference's ``state_dict`` layouts except ``fc.weight`` which is kept in
HWC-flatten order internally (``fcw_permute`` converts for checkpoints).
"""
from __future__ import annotations

import ctypes
from ctypes import c_double, c_float, c_int, c_size_t, c_void_p

import torch

from .. import _native

LAUNCHES = 0


def _count(n: int) -> None:
    global LAUNCHES
    LAUNCHES += n


def _p(t) -> c_void_p:
    return c_void_p(0 if t is None else t.data_ptr())


def _stream() -> c_void_p:
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def _fn(name: str, act):
    """Entry point `name` for the storage dtype of activation tensor `act` (bf16 -> name, fp32 -> name_f32)."""
    lib = _lib()
    if act.dtype == torch.float32:
        return getattr(lib, name + "_f32")
    if act.dtype != torch.bfloat16:
        raise TypeError(f"{name}: activations must be bfloat16 or float32, got {act.dtype}")
    return getattr(lib, name)


def _lib():
    lib = _native.cuda()
    if not getattr(lib, "_b200_ops_bound", False):
        lib.b200_fc_fwd.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_size_t, c_void_p]
        lib.b200_fc_bwd.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_size_t,
                                    c_float, c_int, c_float, c_void_p, c_void_p, c_void_p]
        lib.b200_conv2_dgrad_stats_tc.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                                  c_int, c_int, ctypes.POINTER(c_int), c_void_p]
        lib.b200_sgd.argtypes = [c_void_p, c_void_p, c_size_t, c_float, c_void_p]
        lib.b200_bn_finalize.argtypes = [c_void_p, c_int, c_int, c_double, c_void_p, c_void_p, c_void_p, c_void_p,
                                         c_void_p, c_float, c_float, c_void_p, c_void_p]
        lib.b200_bn_bwd_finalize.argtypes = [c_void_p, c_int, c_int, c_double, c_void_p, c_void_p, c_void_p,
                                             c_float, c_void_p]
        lib.b200_conv1_wgrad.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                         c_int, c_int, c_int, c_float, c_void_p]
        lib.b200_conv1_wgrad_tc.argtypes = lib.b200_conv1_wgrad.argtypes
        lib.b200_conv1_fwd_tc.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                          ctypes.POINTER(c_int), c_void_p]
        lib.b200_conv2_wgrad_tc.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float,
                                            c_void_p]
        lib.b200_conv2_fwd_tc.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                          ctypes.POINTER(c_int), c_void_p]
        lib.b200_conv2_dgrad_tc.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]
        lib.b200_cross_entropy.argtypes = [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_float,
                                           c_void_p]
        lib.b200_reduce_partials.argtypes = [c_void_p, c_int, c_int, c_void_p, c_float, c_void_p]
        lib.b200_l2_flush.argtypes = [c_void_p, c_size_t, c_void_p]
        for name in ("fc_fwd", "fc_bwd", "conv2_dgrad_stats_tc", "conv1_wgrad", "conv1_wgrad_tc", "conv1_fwd_tc", "conv2_wgrad_tc",
                     "conv2_fwd_tc", "conv2_dgrad_tc"):
            getattr(lib, f"b200_{name}_f32").argtypes = getattr(lib, f"b200_{name}").argtypes
        lib._b200_ops_bound = True
    return lib


def sm_count() -> int:
    return int(_lib().b200_sm_count())


def set_reserved_sms(n: int) -> None:
    """SMs the persistent tcgen05 kernels leave to an overlapped collective (parallel/engine_sync.py); 0 = none."""
    _native.check(_lib().b200_set_reserved_sms(int(n)), "set_reserved_sms")


# ---- grid sizes (callers size their partial buffers with these) ------------
def conv1_fwd_grid() -> int:
    return int(_lib().b200_conv1_fwd_grid())


def conv1_wgrad_grid() -> int:
    return int(_lib().b200_conv1_wgrad_grid())


def conv_tc_max_grid() -> int:
    return int(_lib().b200_conv_tc_max_grid())


def pool_bn_bwd_grid() -> int:
    return int(_lib().b200_pool_bn_bwd_grid())


def fc_grid() -> int:
    return int(_lib().b200_fc_grid())


def conv_tc_set_version(v: int) -> None:
    """tcgen05 conv variant (bf16 kernels): 1 = tap-stationary + TMA, 2 = row-stationary + TMA, 3 = row-stationary +
    cp.async, 4 = 3 with the dx-replicated M=128 weight-gradient kernel, 5 = 4 with forward / dgrad fed by one TMA
    request per input row into swizzled pixel rows (default; fp32 kernels support 4 and 5)."""
    _native.check(_lib().b200_conv_tc_set_version(int(v)), "conv_tc_set_version")


def conv_tc_get_version() -> int:
    return int(_lib().b200_conv_tc_get_version())


# ---- kernels ----------------------------------------------------------------
def resize_pil_bilinear(src_u8, xtab, ytab, dst_f32):
    """uint8 [B,h,w] -> fp32 [B,H,W]; bit-exact PIL bilinear + ToTensor (K1/K2)."""
    B, sh, sw = src_u8.shape
    H, W = dst_f32.shape[-2:]
    _native.check(_lib().b200_resize_pil_bilinear(_p(src_u8), B, sh, sw, _p(xtab), _p(ytab), _p(dst_f32), H, W,
                                                  _stream()), "resize_pil_bilinear")
    _count(1)


def conv1_fwd(x_f32, w1, b1, x1_bf16, partials):
    """K3: conv(1->16, 5x5, pad 2) + BN-stat partials. x [B,H,W] fp32 -> x1 [B,H,W,16] bf16."""
    B, H, W = x_f32.shape
    _native.check(_lib().b200_conv1_fwd(_p(x_f32), _p(w1), _p(b1), _p(x1_bf16), _p(partials), B, H, W, _stream()),
                  "conv1_fwd")
    _count(1)


def conv1_fwd_tc(x_f32, w1, b1, x1_bf16, partials) -> int:
    """K3 on tcgen05 (Toeplitz GEMM over 8 output columns, bf16 hi/lo split, fp32 accumulate).
    Returns the number of BN-stat partial rows written (partials: [sm_count()][32])."""
    B, H, W = x_f32.shape
    n = c_int(0)
    _native.check(_fn("b200_conv1_fwd_tc", x1_bf16)(_p(x_f32), _p(w1), _p(b1), _p(x1_bf16), _p(partials), B, H, W,
                                                    ctypes.byref(n), _stream()), "conv1_fwd_tc")
    _count(1)
    return n.value


def conv1_stats_tc(x_f32, w1, b1, partials, row_operand=False) -> int:
    """First pass of the two-pass fp32 layer-1 forward: BatchNorm batch statistics of conv1's output, nothing stored.
    Returns the number of partial rows.  ``row_operand``: the kind::tf32 kernel whose A operand is the raw image row
    (csrc/conv1_hz.cu; W % 4 == 0) instead of the bf16 hi/lo vertical-Toeplitz kernel - use the same choice in pass 2."""
    B, H, W = x_f32.shape
    n = c_int(0)
    lib = _lib()
    if row_operand:
        lib.b200_conv1_stats_tf32.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                              ctypes.POINTER(c_int), c_void_p]
        _native.check(lib.b200_conv1_stats_tf32(_p(x_f32), _p(w1), _p(b1), _p(partials), B, H, W, ctypes.byref(n),
                                                _stream()), "conv1_stats_tf32")
        _count(1)
        return n.value
    lib.b200_conv1_stats_tc_f32.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                            ctypes.POINTER(c_int), c_void_p]
    _native.check(lib.b200_conv1_stats_tc_f32(_p(x_f32), _p(w1), _p(b1), _p(partials), B, H, W, ctypes.byref(n), _stream()),
                  "conv1_stats_tc")
    _count(1)
    return n.value


def conv1_fwd_pool_tc(x_f32, w1, b1, bn, x1_f32, p1_f32, partials, amask_u8=None, row_operand=False):
    """Second pass: conv1 again (bit-identical values) with BN-apply + ReLU + 2x2 max-pool fused into the epilogue ->
    ``p1_f32``; replaces conv1_fwd_tc + bn_relu_pool (K3 + K5).  The conv output goes to ``x1_f32`` for the backward -
    or nowhere (``x1_f32=None``) when ``amask_u8`` ([B,H/2,W/2,16] uint8: arg-max | ReLU << 2) is produced for
    ``conv1_wgrad_nox1``."""
    B, H, W = x_f32.shape
    assert p1_f32.dtype == torch.float32 and (x1_f32 is None or x1_f32.dtype == torch.float32)
    lib = _lib()
    if row_operand:
        assert x1_f32 is None and amask_u8 is not None, "the row-operand kernel stores no conv output"
        assert amask_u8.dtype == torch.uint8 and amask_u8.numel() == p1_f32.numel()
        lib.b200_conv1_fwd_pool_mask_tf32.argtypes = [c_void_p] * 6 + [c_int, c_int, c_int, c_void_p]
        _native.check(lib.b200_conv1_fwd_pool_mask_tf32(_p(x_f32), _p(w1), _p(b1), _p(bn), _p(p1_f32), _p(amask_u8),
                                                        B, H, W, _stream()), "conv1_fwd_pool_mask_tf32")
        _count(1)
        return
    if amask_u8 is not None:
        assert amask_u8.dtype == torch.uint8 and amask_u8.numel() == p1_f32.numel()
        lib.b200_conv1_fwd_pool_mask_tc_f32.argtypes = [c_void_p] * 8 + [c_int, c_int, c_int, c_void_p]
        _native.check(lib.b200_conv1_fwd_pool_mask_tc_f32(_p(x_f32), _p(w1), _p(b1), _p(bn), _p(x1_f32), _p(p1_f32),
                                                          _p(amask_u8), _p(partials), B, H, W, _stream()),
                      "conv1_fwd_pool_mask_tc")
        _count(1)
        return
    assert x1_f32 is not None
    lib.b200_conv1_fwd_pool_tc_f32.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                               c_int, c_int, c_int, c_void_p]
    _native.check(lib.b200_conv1_fwd_pool_tc_f32(_p(x_f32), _p(w1), _p(b1), _p(bn), _p(x1_f32), _p(p1_f32), _p(partials),
                                                 B, H, W, _stream()), "conv1_fwd_pool_tc")
    _count(1)


def conv1_wgrad_nox1_partial() -> int:
    return int(_lib().b200_conv1_wgrad_nox1_partial())


def conv1_wgrad_nox1(x_f32, dp1_f32, amask_u8, w1, b1, bn, red, partials, dw1, gscale=1.0):
    """K13 without the conv1 output: routes dp1 through the forward's arg-max/ReLU bytes on the tensor cores and gets the
    dense BatchNorm term from the input's 25x25 Gram matrix (same kernel, one more MMA per K-step).  fp32 storage,
    H % 8 == 0.  partials: [conv1_wgrad_grid()][conv1_wgrad_nox1_partial()]."""
    B, H, W = x_f32.shape
    lib = _lib()
    lib.b200_conv1_wgrad_nox1_f32.argtypes = [c_void_p] * 9 + [c_int, c_int, c_int, c_float, c_void_p]
    _native.check(lib.b200_conv1_wgrad_nox1_f32(_p(x_f32), _p(dp1_f32), _p(amask_u8), _p(w1), _p(b1), _p(bn), _p(red),
                                                _p(partials), _p(dw1), B, H, W, gscale, _stream()), "conv1_wgrad_nox1")
    _count(2)


def bn_finalize(partials, nparts, C, count, gamma, beta, running_mean, running_var, nbt, out, momentum=0.1,
                eps=1e-5):
    """K4: batch stats -> [scale|shift|mean|rstd] + running-stat update."""
    _native.check(_lib().b200_bn_finalize(_p(partials), nparts, C, float(count), _p(gamma), _p(beta),
                                          _p(running_mean), _p(running_var), _p(nbt), momentum, eps, _p(out),
                                          _stream()), "bn_finalize")
    _count(1)


def bn_relu_pool(x_bf16, bn, p_bf16):
    """K5/K7: BN-apply + ReLU + 2x2 max-pool, bf16 NHWC -> bf16 NHWC."""
    B, H, W, C = x_bf16.shape
    assert x_bf16.dtype == p_bf16.dtype
    _native.check(_fn("b200_bn_relu_pool", x_bf16)(_p(x_bf16), _p(bn), _p(p_bf16), B, H, W, C, _stream()),
                  "bn_relu_pool")
    _count(1)


def pool_bn_bwd(x_bf16, dp_bf16, bn, red, partials, dx_bf16, mode):
    """K11/K13: mode 0 = reduction partials, mode 1 = dense dx (may alias x)."""
    B, H, W, C = x_bf16.shape
    assert x_bf16.dtype == dp_bf16.dtype and (dx_bf16 is None or dx_bf16.dtype == x_bf16.dtype)
    _native.check(_fn("b200_pool_bn_bwd", x_bf16)(_p(x_bf16), _p(dp_bf16), _p(bn), _p(red), _p(partials),
                                                  _p(dx_bf16), B, H, W, C, mode, _stream()), "pool_bn_bwd")
    _count(1)


def bn_bwd_finalize(partials, nparts, C, count, red, dgamma, dbeta, gscale=1.0):
    _native.check(_lib().b200_bn_bwd_finalize(_p(partials), nparts, C, float(count), _p(red), _p(dgamma),
                                              _p(dbeta), gscale, _stream()), "bn_bwd_finalize")
    _count(1)


def conv1_wgrad(x_f32, x1_bf16, dp1_bf16, bn, red, partials, dw1, gscale=1.0):
    """K13: fused pool1/ReLU1/BN1 backward + conv1 weight gradient."""
    B, H, W = x_f32.shape
    assert x1_bf16.dtype == dp1_bf16.dtype
    _native.check(_fn("b200_conv1_wgrad", x1_bf16)(_p(x_f32), _p(x1_bf16), _p(dp1_bf16), _p(bn), _p(red),
                                                   _p(partials), _p(dw1), B, H, W, gscale, _stream()), "conv1_wgrad")
    _count(2)


def conv1_wgrad_tc(x_f32, x1_bf16, dp1_bf16, bn, red, partials, dw1, gscale=1.0):
    """K13 on tcgen05: same contract as conv1_wgrad; pixels are the GEMM K dimension, the conv1-output
    gradient is produced into shared memory by CUDA-core warps and never touches HBM."""
    B, H, W = x_f32.shape
    assert x1_bf16.dtype == dp1_bf16.dtype
    _native.check(_fn("b200_conv1_wgrad_tc", x1_bf16)(_p(x_f32), _p(x1_bf16), _p(dp1_bf16), _p(bn), _p(red),
                                                      _p(partials), _p(dw1), B, H, W, gscale, _stream()),
                  "conv1_wgrad_tc")
    _count(2)


def conv2_fwd_tc(p1_bf16, w2, b2, x2_bf16, partials) -> int:
    """K6 (tcgen05): conv(16->32) + BN-stat partials. Returns the number of partial rows."""
    B, H, W, _ = p1_bf16.shape
    n = c_int(0)
    assert p1_bf16.dtype == x2_bf16.dtype
    _native.check(_fn("b200_conv2_fwd_tc", p1_bf16)(_p(p1_bf16), _p(w2), _p(b2), _p(x2_bf16), _p(partials), B, H, W,
                                                    ctypes.byref(n), _stream()), "conv2_fwd_tc")
    _count(1)
    return n.value


def conv2_dgrad_tc(dx2_bf16, w2, dp1_bf16):
    """K12a (tcgen05): data gradient of conv2, [B,H,W,32] -> [B,H,W,16]."""
    B, H, W, _ = dx2_bf16.shape
    assert dx2_bf16.dtype == dp1_bf16.dtype
    _native.check(_fn("b200_conv2_dgrad_tc", dx2_bf16)(_p(dx2_bf16), _p(w2), _p(dp1_bf16), B, H, W, _stream()),
                  "conv2_dgrad_tc")
    _count(1)


def conv2_wgrad_tc(p1_bf16, dx2_bf16, partials, dw2, gscale=1.0):
    """K12b (tcgen05): weight gradient of conv2 -> [32,16,5,5] fp32."""
    B, H, W, _ = p1_bf16.shape
    assert p1_bf16.dtype == dx2_bf16.dtype
    _native.check(_fn("b200_conv2_wgrad_tc", p1_bf16)(_p(p1_bf16), _p(dx2_bf16), _p(partials), _p(dw2), B, H, W,
                                                      gscale, _stream()), "conv2_wgrad_tc")
    _count(2)


def fc_fwd(p2_bf16, W, bias, partials, logits):
    """K8: logits for a batch chunk of <= 8 rows. p2 [bt,F] bf16, W [10,F] fp32."""
    bt, F = p2_bf16.shape[0], p2_bf16[0].numel()
    _native.check(_fn("b200_fc_fwd", p2_bf16)(_p(p2_bf16), _p(W), _p(bias), _p(partials), _p(logits), bt, F,
                                              _stream()), "fc_fwd")
    _count(2)


def cross_entropy(logits, labels, loss, dlogits, dbias, gscale=1.0):
    """K9: mean CE loss + dlogits + dbias in one block."""
    _native.check(_lib().b200_cross_entropy(_p(logits), _p(labels), logits.shape[0], _p(loss), _p(dlogits),
                                            _p(dbias), gscale, _stream()), "cross_entropy")
    _count(1)


def fc_bwd(dlogits, p2_bf16, W, dW, dp2_bf16, lr=0.0, accumulate=False, gscale=1.0, bn2=None, bwd_partials=None):
    """K10(+K14): dW (optional), dp2 and optional in-place SGD on W, one streaming pass over W for ANY batch (up to 8
    rows are held in registers; larger batches loop inside the pass).  With ``bn2`` the BatchNorm-2 backward
    reductions are produced too (partials [fc_grid()][64]), replacing a pass over x2."""
    bt, F = p2_bf16.shape[0], p2_bf16[0].numel()
    assert p2_bf16.dtype == dp2_bf16.dtype
    # dW mode of the C entry point: 0 = overwrite fp32, 1 = accumulate fp32, 2 = overwrite a bf16 bucket (DDP gradient
    # exchange in bf16)
    mode = 1 if accumulate else 0
    if dW is not None and dW.dtype == torch.bfloat16:
        assert not accumulate, "a bf16 gradient bucket is overwritten, never accumulated into"
        mode = 2
    _native.check(_fn("b200_fc_bwd", p2_bf16)(_p(dlogits), _p(p2_bf16), _p(W), _p(dW), _p(dp2_bf16), bt, F, lr,
                                              mode, gscale, _p(bn2), _p(bwd_partials), _stream()),
                  "fc_bwd")
    _count(1)


def fc_bwd_tile_slab_floats(sweeps_per_slab: int) -> int:
    lib = _lib()
    lib.b200_fc_bwd_tile_slab_floats.restype = c_size_t
    return int(lib.b200_fc_bwd_tile_slab_floats(int(sweeps_per_slab)))


def fc_bwd_tiles(dlogits, p2, W, dW, dp2, ready_u32, sweeps_per_slab, gscale=1.0, bn2=None, bwd_partials=None):
    """K10 for data-parallel runs: fc backward that bumps ``ready_u32[t]`` (symmetric memory) once per block when
    its share of column slab t of dW is written, so the tile-driven bucket all-reduce (K16) running beside it can
    start on a slab while the rest of dW is still being produced.  bt <= 8."""
    lib = _lib()
    lib.b200_fc_bwd_tiles.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_size_t, c_float,
                                      c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]
    bt, F = p2.shape[0], p2[0].numel()
    assert p2.dtype == dp2.dtype and ready_u32.dtype in (torch.int32, torch.uint32)
    _native.check(lib.b200_fc_bwd_tiles(_p(dlogits), _p(p2), _p(W), _p(dW), _p(dp2), bt, F, gscale, _p(bn2),
                                        _p(bwd_partials), _p(ready_u32), int(sweeps_per_slab),
                                        1 if p2.dtype == torch.float32 else 0, _stream()), "fc_bwd_tiles")
    _count(1)


def conv2_dgrad_stats_tc(dx2_bf16, w2, pooled_p1, bn1, dp1_bf16, partials) -> int:
    """K12a + K13-stats: conv2 data gradient with the BatchNorm-1 backward reductions fused into its epilogue
    (``pooled_p1`` may alias ``dp1_bf16``).  Returns the number of partial rows ([rows][32])."""
    B, H, W, _ = dx2_bf16.shape
    n = c_int(0)
    assert dx2_bf16.dtype == dp1_bf16.dtype == pooled_p1.dtype
    _native.check(_fn("b200_conv2_dgrad_stats_tc", dx2_bf16)(_p(dx2_bf16), _p(w2), _p(pooled_p1), _p(bn1),
                                                             _p(dp1_bf16), _p(partials), B, H, W, ctypes.byref(n),
                                                             _stream()), "conv2_dgrad_stats_tc")
    _count(1)
    return n.value


def sgd(p, g, lr):
    """K14: p -= lr * g."""
    _native.check(_lib().b200_sgd(_p(p), _p(g), p.numel(), lr, _stream()), "sgd")
    _count(1)


def fcw_permute(src, dst, C, Hd, Wd, to_hwc: bool):
    _native.check(_lib().b200_fcw_permute(_p(src), _p(dst), C, Hd, Wd, 1 if to_hwc else 0, _stream()),
                  "fcw_permute")
    _count(1)


def nhwc_bf16_to_nchw_f32(src, dst):
    B, H, W, C = src.shape
    _native.check(_lib().b200_nhwc_bf16_to_nchw_f32(_p(src), _p(dst), B, C, H, W, _stream()), "nhwc->nchw")
    _count(1)


def nchw_f32_to_nhwc_bf16(src, dst):
    B, C, H, W = src.shape
    _native.check(_lib().b200_nchw_f32_to_nhwc_bf16(_p(src), _p(dst), B, C, H, W, _stream()), "nchw->nhwc")
    _count(1)


def l2_flush(buf):
    _native.check(_lib().b200_l2_flush(_p(buf), buf.numel() * buf.element_size(), _stream()), "l2_flush")
