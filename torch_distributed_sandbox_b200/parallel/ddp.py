"""DistributedDataParallel - the data-parallel engine (L4 / D7 in SURVEY.md).

Reference use: ``nn.parallel.DistributedDataParallel(model, device_ids=[gpu])`` with every
other argument at its default (mnist_distributed.py:67).  What those defaults do inside
torch's C++ Reducer, and what this class therefore provides itself:

* constructor: parameter shapes are verified across ranks, then parameters and buffers are
  broadcast from rank 0 (``init_sync``);
* every forward: buffers (the BatchNorm running stats) are re-broadcast from rank 0
  (``broadcast_buffers=True``);
* backward: per-parameter post-accumulate hooks copy each gradient, pre-divided by the
  world size, into a flat bucket; a full bucket is all-reduced (SUM) asynchronously while
  backward continues; at the end of backward the buckets are waited on and copied back
  into ``.grad``.  Iteration 1 uses a single bucket; afterwards buckets are rebuilt in
  gradient-arrival order with caps 1 MiB (first) / 25 MiB (rest);
* ``state_dict()`` keys carry the ``module.`` prefix.

Two wrapped-module kinds are supported:
  - any ``torch.nn.Module`` (generic path above; CPU tensors on the host backend, CUDA
    tensors on the native backend with buckets living in symmetric memory), and
  - the fused engine (``models.FusedModule``): its fc backward writes the 720 MB
    ``fc.weight`` gradient straight into a symmetric-memory bucket and the all-reduce kernel
    runs on a side stream under the conv backward (``EngineGradSync``, K10 -> K16).
"""
from __future__ import annotations

import contextlib
import hashlib
from typing import List, Optional

import torch
from torch import nn

from .. import distributed as dist
from .buckets import DEFAULT_BUCKET_CAP_MB, FIRST_BUCKET_BYTES, bucket_offsets, plan_buckets


class _Bucket:
    def __init__(self, params: List[nn.Parameter], backend, world: int):
        self.params = params
        numels = [p.numel() for p in params]
        self.offsets, total = bucket_offsets(numels, 4)
        p0 = params[0]
        self.numel = total
        # On the native backend the flat buffer is symmetric memory (peer-mapped), so the
        # all-reduce kernel can read every rank's copy directly; elsewhere a plain tensor.
        alloc = getattr(backend, "alloc_symmetric", None)
        self.flat = alloc(total, p0.dtype) if (alloc is not None and p0.is_cuda) else \
            torch.zeros(total, dtype=p0.dtype, device=p0.device)
        self.views = [self.flat[o:o + n].view(p.shape) for o, n, p in zip(self.offsets, numels, params)]
        self.pending = len(params)
        self.work = None

    def bytes(self) -> int:
        return sum(p.numel() * p.element_size() for p in self.params)


class DistributedDataParallel(nn.Module):
    _instances = 0

    def __init__(self, module: nn.Module, device_ids=None, output_device=None, dim: int = 0,
                 broadcast_buffers: bool = True, init_sync: bool = True, process_group=None,
                 bucket_cap_mb: Optional[float] = None, find_unused_parameters: bool = False,
                 gradient_as_bucket_view: bool = False, static_graph: bool = False,
                 gradient_dtype: Optional[str] = None):
        super().__init__()
        if not dist.is_initialized():
            raise RuntimeError("DistributedDataParallel needs an initialised process group")
        if find_unused_parameters:
            raise NotImplementedError("find_unused_parameters=True is not supported (the reference uses the default False)")
        self.module = module
        self.process_group = process_group or dist._default()
        self.world_size = self.process_group.size()
        self.broadcast_buffers = broadcast_buffers
        self.bucket_bytes_cap = int((bucket_cap_mb if bucket_cap_mb is not None else DEFAULT_BUCKET_CAP_MB) * 1024 * 1024)
        self.device_ids = device_ids
        self.require_backward_grad_sync = True
        self._engine_sync = None
        engine = getattr(module, "engine", None)
        if engine is not None:
            # fused engine: gradient exchange is wired into the engine's own backward
            from .engine_sync import EngineGradSync
            self._verify_shapes([("fc.weight", engine.fcw)] + list(engine.params.items()))
            self._engine_sync = EngineGradSync(engine, self.process_group, init_sync=init_sync,
                                               broadcast_buffers=broadcast_buffers, gradient_dtype=gradient_dtype)
            return
        self._params = [p for p in module.parameters() if p.requires_grad]
        if not self._params:
            raise RuntimeError("DistributedDataParallel is not needed when a module doesn't have any parameter that requires a gradient")
        self._verify_shapes(list(module.named_parameters()))
        if init_sync:
            self._sync_module_states()
        # iteration 1: a single bucket over all parameters (torch passes [sys.maxsize] when
        # find_unused_parameters is False), in reverse registration order
        self._buckets: List[_Bucket] = []
        self._build_buckets(list(reversed(self._params)), [1 << 62])
        self._rebuilt = False
        self._arrival: List[nn.Parameter] = []
        self._callback_queued = False
        self._hooks = [p.register_post_accumulate_grad_hook(self._make_hook(p)) for p in self._params]

    # ------------------------------------------------------------- bring-up --
    def _verify_shapes(self, named) -> None:
        """C3: shapes must agree on every rank (torch all-gathers the metadata; a digest through
        the store is enough on this cold path)."""
        desc = ";".join(f"{n}:{tuple(p.shape)}:{p.dtype}" for n, p in named)
        digest = hashlib.sha1(desc.encode()).hexdigest()
        store, g = dist.get_store(), self.process_group
        DistributedDataParallel._instances += 1           # same construction order on every rank
        key = f"ddp_verify/{g.id}/{DistributedDataParallel._instances}"
        store.set(f"{key}/{g.rank()}", digest)
        ref = store.get(f"{key}/0").decode()
        if ref != digest:
            raise RuntimeError(f"DDP: parameter shapes/dtypes on rank {dist.get_rank()} differ from rank {g.ranks[0]}")

    def _sync_module_states(self) -> None:
        """C2: parameters and buffers take rank 0's values."""
        src = self.process_group.ranks[0]
        with torch.no_grad():
            for t in list(self.module.parameters()) + list(self.module.buffers()):
                dist.broadcast(t.data, src=src, group=self.process_group)

    def _sync_buffers(self) -> None:
        """C4: per-forward buffer broadcast from rank 0."""
        src = self.process_group.ranks[0]
        with torch.no_grad():
            for b in self.module.buffers():
                dist.broadcast(b.data, src=src, group=self.process_group)

    # -------------------------------------------------------------- buckets --
    def _build_buckets(self, ordered: List[nn.Parameter], limits) -> None:
        plan = plan_buckets([p.data for p in ordered], limits)
        # the iteration-1 bucket spans the whole model (720 MB for the 3000x3000 net): give its symmetric memory
        # back before the rebuilt buckets are allocated instead of holding both until the group is destroyed
        free = getattr(self.process_group.backend, "free_symmetric", None)
        old, self._buckets, self._where = getattr(self, "_buckets", []), [], {}
        for b in old:
            flat, b.flat, b.views = b.flat, None, []
            if free is not None and flat is not None and flat.is_cuda:
                free(flat)
        self._buckets = [_Bucket([ordered[i] for i in idx], self.process_group.backend, self.world_size)
                         for idx in plan]
        self._where = {}
        for b in self._buckets:
            for j, p in enumerate(b.params):
                self._where[p] = (b, j)

    def bucket_sizes(self) -> List[int]:
        """Bytes per gradient bucket, in launch order (tests compare with torch's planner)."""
        if self._engine_sync is not None:
            return self._engine_sync.bucket_sizes()
        return [b.bytes() for b in self._buckets]

    def _make_hook(self, p):
        def hook(param):
            if not self.require_backward_grad_sync:
                return
            if not self._callback_queued:
                self._callback_queued = True
                torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
            if not self._rebuilt:
                self._arrival.append(param)
            b, j = self._where[param]
            b.views[j].copy_(param.grad)
            b.views[j].div_(self.world_size)          # pre-divide, then SUM (reducer div_factor)
            b.pending -= 1
            if b.pending == 0:
                b.work = dist.all_reduce(b.flat, dist.ReduceOp.SUM, group=self.process_group, async_op=True)
        return hook

    def _finalize_backward(self) -> None:
        self._callback_queued = False
        for b in self._buckets:
            if b.pending != 0:
                raise RuntimeError("DDP: backward finished but a gradient bucket is incomplete - a parameter "
                                   "did not receive a gradient (find_unused_parameters is not supported)")
            if b.work is not None:
                b.work.wait()
            for p, v in zip(b.params, b.views):
                p.grad.copy_(v)
            b.pending = len(b.params)
            b.work = None
        if not self._rebuilt:
            # rebuild once, in the order gradients actually arrived (reducer rebuild_buckets)
            self._rebuilt = True
            if len(self._arrival) == len(self._params):
                self._build_buckets(self._arrival, [FIRST_BUCKET_BYTES, self.bucket_bytes_cap])
            self._arrival = []

    # -------------------------------------------------------------- forward --
    def forward(self, *args, **kwargs):
        if self._engine_sync is not None:
            self._engine_sync.before_forward()
            return self.module(*args, **kwargs)
        if self.broadcast_buffers and self.require_backward_grad_sync and any(True for _ in self.module.buffers()):
            self._sync_buffers()
        return self.module(*args, **kwargs)

    def train_step(self, images, labels):
        """Fused fast path of an engine-backed module (forward + loss + backward + all-reduce +
        SGD); includes the per-forward buffer broadcast DDP performs."""
        if self._engine_sync is None:
            raise AttributeError("train_step is only available when wrapping a FusedModule")
        self._engine_sync.before_forward()
        return self.module.train_step(images, labels)

    @contextlib.contextmanager
    def no_sync(self):
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old
