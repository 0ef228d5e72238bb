"""Parallelism strategies.  The reference has exactly one - data parallel -
and so does this package: DDP wrapper, gradient bucketing, shard sampler, and the
symmetric-memory collectives they run on."""
from .sampler import DistributedSampler  # noqa: F401
from .buckets import plan_buckets, FIRST_BUCKET_BYTES, DEFAULT_BUCKET_CAP_MB  # noqa: F401


def __getattr__(name):
    if name == "DistributedDataParallel":
        from .ddp import DistributedDataParallel
        return DistributedDataParallel
    raise AttributeError(name)
