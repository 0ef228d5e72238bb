"""Gradient bucket planning (Python face of csrc/reducer.cpp; D7 in SURVEY.md).

torch's DDP, which the reference instantiates with default arguments at
mnist_distributed.py:67, groups gradients into flat buckets: first cap 1 MiB, then 25 MiB,
one big bucket on the very first iteration, rebuilt in gradient-arrival order afterwards.
"""
from __future__ import annotations

import ctypes
from ctypes import c_int, c_int32, c_int64
from typing import List, Sequence

import torch

from .. import _native

FIRST_BUCKET_BYTES = 1 << 20          # torch: kDefaultFirstBucketBytes
DEFAULT_BUCKET_CAP_MB = 25            # torch: kDefaultBucketBytesCap = 25 MiB
_DT_KEYS = {}


def _key(t: torch.Tensor) -> int:
    k = (t.dtype, str(t.device))
    return _DT_KEYS.setdefault(k, len(_DT_KEYS))


def plan_buckets(tensors: Sequence[torch.Tensor], limits: Sequence[int]) -> List[List[int]]:
    """Indices of ``tensors`` grouped into buckets, in closing order."""
    n = len(tensors)
    if n == 0:
        return []
    sizes = (c_int64 * n)(*[t.numel() * t.element_size() for t in tensors])
    keys = (c_int32 * n)(*[_key(t) for t in tensors])
    lim = (c_int64 * len(limits))(*[int(x) for x in limits])
    out = (c_int32 * n)()
    lib = _native.host()
    lib.b200h_plan_buckets.restype = c_int
    nb = lib.b200h_plan_buckets(sizes, keys, n, lim, len(limits), out)
    buckets: List[List[int]] = [[] for _ in range(nb)]
    for i in range(n):
        buckets[out[i]].append(i)
    return buckets


def bucket_offsets(numels: Sequence[int], align_elems: int = 4):
    n = len(numels)
    arr = (c_int64 * n)(*[int(x) for x in numels])
    offs = (c_int64 * n)()
    lib = _native.host()
    lib.b200h_bucket_offsets.restype = c_int64
    total = lib.b200h_bucket_offsets(arr, n, align_elems, offs)
    return list(offs), int(total)
