"""Host-side coefficient tables for the GPU resize kernel (K2).

The reference input pipeline is ``Resize([3000, 3000])`` + ``ToTensor`` on a PIL
``L`` image (mnist_onegpu.py:53, mnist_distributed.py:71): Pillow's 8-bit
bilinear resample - horizontal pass, then vertical pass, 22-bit fixed-point
weights, each pass rounded and clipped to uint8 - followed by ``/255``.  The
tables below restate Pillow's coefficient pre-computation for the triangle
filter so that the device kernel (and the numpy emulation used as a CPU oracle)
reproduce the reference's pixels bit for bit.  Only up-scaling (or identity) is
supported: with ``in <= out`` every output pixel has at most two taps.
"""
from __future__ import annotations

import functools

import numpy as np

_PRECISION_BITS = 32 - 8 - 2


@functools.lru_cache(maxsize=32)
def pil_bilinear_table(in_size: int, out_size: int) -> np.ndarray:
    """int32 [out_size, 3] rows of (first source index, weight0, weight1)."""
    if in_size > out_size:
        raise ValueError("GPU resize supports up-scaling only (the reference goes 28 -> 3000)")
    scale = in_size / out_size
    support = 1.0  # bilinear support * max(scale, 1)
    centers = (np.arange(out_size, dtype=np.float64) + 0.5) * scale
    xmin = np.trunc(centers - support + 0.5).astype(np.int64)
    xmin = np.maximum(xmin, 0)
    xmax = np.trunc(centers + support + 0.5).astype(np.int64)
    xmax = np.minimum(xmax, in_size)
    ntap = xmax - xmin
    assert ntap.max() <= 2 and ntap.min() >= 1
    w = np.zeros((out_size, 2), dtype=np.float64)
    for t in range(2):
        arg = np.abs(t + xmin - centers + 0.5)
        wt = np.where(arg < 1.0, 1.0 - arg, 0.0)
        w[:, t] = np.where(t < ntap, wt, 0.0)
    w /= w.sum(axis=1, keepdims=True)
    scaled = w * (1 << _PRECISION_BITS)
    k = np.where(scaled < 0, np.trunc(scaled - 0.5), np.trunc(scaled + 0.5)).astype(np.int64)
    tab = np.empty((out_size, 3), dtype=np.int32)
    tab[:, 0] = xmin
    tab[:, 1] = k[:, 0]
    tab[:, 2] = k[:, 1]
    return tab


def resize_u8_reference(src: np.ndarray, out_hw) -> np.ndarray:
    """numpy emulation of the device kernel: uint8 [..., h, w] -> uint8 [..., H, W]."""
    h, w = src.shape[-2:]
    H, W = out_hw
    xt, yt = pil_bilinear_table(w, W), pil_bilinear_table(h, H)
    half = 1 << (_PRECISION_BITS - 1)
    s = src.astype(np.int64)
    x0, x1 = xt[:, 0], np.minimum(xt[:, 0] + 1, w - 1)
    hor = (s[..., :, x0] * xt[:, 1] + s[..., :, x1] * xt[:, 2] + half) >> _PRECISION_BITS
    hor = np.clip(hor, 0, 255)
    y0, y1 = yt[:, 0], np.minimum(yt[:, 0] + 1, h - 1)
    ver = (hor[..., y0, :] * yt[:, 1][:, None] + hor[..., y1, :] * yt[:, 2][:, None] + half) >> _PRECISION_BITS
    return np.clip(ver, 0, 255).astype(np.uint8)
