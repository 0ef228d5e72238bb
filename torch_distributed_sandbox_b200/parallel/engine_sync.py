"""Gradient synchronisation of the fused engine under data parallelism (D7 / C1-C4).

What torch's Reducer does for the reference (mnist_distributed.py:67) - bucket the
gradients, pre-divide, all-reduce during backward, copy back, then ``optimizer.step()`` -
is re-cut here around the model's actual shape: one 720 MB gradient (fc.weight, the FIRST
one backward produces) and 53 KB of everything else.

  * fc.weight's gradient is written by the fc backward kernel straight into a symmetric
    bucket (no flatten copy, no copy-back);
  * as soon as that kernel is enqueued, a side stream launches the two-shot all-reduce
    (NVLS multimem when available, P2P otherwise) with the 1/world scale and the SGD update
    fused into its epilogue; it runs under the whole conv backward on a few CTAs;
  * the 53 KB bucket is one one-shot kernel at the end of backward;
  * BatchNorm buffers are re-broadcast from rank 0 before every forward (DDP's
    ``broadcast_buffers=True``) as a single 512-byte kernel.
"""
from __future__ import annotations

import os
from typing import Dict, List

import torch

from .. import distributed as dist
from ..utils import trace


class EngineGradSync:
    def __init__(self, engine, group, init_sync: bool = True, broadcast_buffers: bool = True,
                 gradient_dtype: str = None):
        self.engine, self.group = engine, group
        # "bf16": the fc.weight gradient (99.99 % of the bytes) is produced, exchanged and reduced as bf16 - what torch's
        # bf16_compress_hook does for DDP - with fp32 accumulation in the switch and the fp32 SGD update fused into the
        # same pass.  Lower gradient precision than the reference's fp32 exchange: opt-in.
        self.gradient_dtype = (gradient_dtype or os.environ.get("B200DP_GRAD_DTYPE", "fp32")).lower()
        if self.gradient_dtype not in ("fp32", "bf16"):
            raise ValueError(f"gradient_dtype must be 'fp32' or 'bf16', not {self.gradient_dtype!r}")
        self.backend = group.backend
        self.world = group.size()
        self.broadcast_buffers = broadcast_buffers
        self.fuse_sgd = os.environ.get("B200DP_FUSE_AR_SGD", "1") == "1"
        self.algo = os.environ.get("B200DP_AR_ALGO") or None         # "nvls" | "p2p" | None (auto)
        self.blocks = int(os.environ.get("B200DP_AR_BLOCKS", "32"))     # CTAs of the overlapped bucket all-reduce (tuned at N=2)
        self._fc_event = None
        self._steps = 0
        self._bn_synced = False
        # SM partition (common.cu); 0 = share SMs.  Each rank reduces 1/world of the bucket, so the comm CTAs a
        # step needs shrink with the world size (measured best: 8 at N=2, 4 at N=8).
        self.reserved_sms = int(os.environ.get("B200DP_RESERVED_SMS", str({2: 8, 4: 6}.get(self.world, 4))))
        # When the fc.weight bucket all-reduce starts:
        #   "tile"  with the fc backward itself: fc_bwd reports finished column slabs of dW through counters in symmetric
        #           memory and the all-reduce kernel, launched beside it, consumes each slab as soon as every rank has
        #           produced it (K10 -> K16 tile pipeline);
        #   "fc"    when the fc backward kernel has finished;  "pool"  one kernel later (after the HBM-bound pool/BN backward).
        self.ar_start = os.environ.get("B200DP_AR_START", "pool")
        self.start_after_pool = self.ar_start == "pool"
        self.tile_sweeps = int(os.environ.get("B200DP_TILE_SWEEPS", "2"))      # fc_bwd grid sweeps per slab (~48 MB of dW)
        self.tile_ready = None
        self._tile_steps = 0
        # "onepass": run the whole-bucket exchange with the one-pass kernel of the tile pipeline (reduce + 1/N + SGD +
        # weight push in a single sweep, no reduce-scatter / all-gather split) instead of the two-shot kernel;
        # with B200DP_AR_START=pool|fc its slab counters are simply not waited for (target 0).
        self.onepass = os.environ.get("B200DP_AR_KERNEL", "twoshot") == "onepass"
        self._onepass_ready = None
        self._debug_skip_fc = os.environ.get("B200DP_DEBUG_SKIP_FC_ALLREDUCE", "0") == "1"   # timing experiments only
        be = self.backend
        if self.world > 1 and getattr(be, "name", "") != "native":
            raise RuntimeError("the fused engine needs the native (symmetric-memory) backend for world_size > 1")
        engine.comm = self
        engine.grad_scale = 1.0            # the 1/world scale is applied inside the all-reduce kernels
        if self.world > 1:
            n = 10 * engine.F
            with torch.cuda.device(engine.device):
                bn = be.alloc_symmetric(128, torch.float32, multicast=False)       # (first: brings the backend up)
                if self.gradient_dtype == "bf16" and not (self.fuse_sgd and be.multicast and self.ar_start != "tile"
                                                          and not self.onepass and n % 8 == 0):
                    import warnings
                    warnings.warn("b200dp: bf16 gradient exchange needs NVLS multicast and the fused SGD epilogue "
                                  "(and neither the tile pipeline nor the one-pass kernel); exchanging fp32")
                    self.gradient_dtype = "fp32"
                grad = be.alloc_symmetric(n, torch.bfloat16 if self.gradient_dtype == "bf16" else torch.float32)
                wsym = be.alloc_symmetric(n, torch.float32) if self.fuse_sgd else None
                sg = be.alloc_symmetric(engine.small_grad.numel(), torch.float32, multicast=False)
                engine.rebind_storage(fcw=wsym, fcw_grad=grad, bn_flat=bn, small_grad=sg)
                if self.ar_start == "tile" and self.fuse_sgd and engine.B <= 8:
                    self.tile_ready = be.alloc_symmetric(256, torch.int32, multicast=False)     # slab counters
                elif self.onepass and self.fuse_sgd:
                    self._onepass_ready = be.alloc_symmetric(256, torch.int32, multicast=False)  # stays zero
            self.comm_stream = be.comm_stream
            if self.reserved_sms > 0 and be.multicast and (self.algo in (None, "nvls", "nvls_excl")):
                from .. import _native
                _native.check(_native.cuda().b200_set_reserved_sms(self.reserved_sms), "set_reserved_sms")
                self.algo, self.blocks = "nvls_excl", self.reserved_sms
            elif self.reserved_sms > 0 and (self.tile_ready is not None or self._onepass_ready is not None):
                from .. import _native                     # the tile kernel's CTAs also own their SMs (P2P loads)
                _native.check(_native.cuda().b200_set_reserved_sms(self.reserved_sms), "set_reserved_sms")
        if self.tile_ready is not None:
            from .. import ops
            self.tile_slab_floats = ops.fc_bwd_tile_slab_floats(self.tile_sweeps)
            self.tile_blocks = ops.fc_grid()
            assert (engine.F + self.tile_slab_floats - 1) // self.tile_slab_floats <= self.tile_ready.numel()
            # Load the producer kernel NOW.  CUDA loads a kernel lazily at its first launch, and that load can wait for
            # running kernels to drain: a first launch of fc_bwd while the consumer is already spinning on its counters
            # would deadlock (seen on the first step: the launch call itself blocked until the consumer timed out).
            with torch.cuda.device(engine.device):
                Bw = min(engine.B, 8)
                dummy_p2 = torch.zeros(Bw, 32, device=engine.device, dtype=engine.act_dtype)
                ops.fc_bwd_tiles(torch.zeros(Bw, 10, device=engine.device), dummy_p2,
                                 torch.zeros(10, 32, device=engine.device), torch.zeros(10, 32, device=engine.device),
                                 torch.empty_like(dummy_p2), torch.zeros(4, dtype=torch.int32, device=engine.device),
                                 self.tile_sweeps)
                torch.cuda.synchronize(engine.device)
        if init_sync:
            self._sync_params()

    # ---------------------------------------------------------------- setup --
    def _sync_params(self) -> None:
        """C2: everything takes rank 0's values (cold path)."""
        e, src = self.engine, self.group.ranks[0]
        for t in (e.small, e.fcw, e.bn_flat):
            dist.broadcast(t, src=src, group=self.group)

    def describe(self) -> Dict:
        d = dict(self.backend.describe()) if hasattr(self.backend, "describe") else {}
        mc = bool(d.get("nvls_multicast"))
        algo = self.algo or ("nvls" if mc else "p2p")
        if self.tile_ready is not None:
            algo = ("nvls" if mc and self.algo != "p2p" else "p2p") + f"-tiles({self.tile_sweeps} sweeps/slab, starts with fc_bwd)"
        elif self._onepass_ready is not None:
            algo = ("nvls" if mc and self.algo not in ("p2p",) else "p2p") + f"(starts after {self.ar_start})"
        d.update({"fc_bucket_algo": (f"one-pass-{algo}" if (self.tile_ready is not None or self._onepass_ready is not None)
                                     else f"two-shot-{algo}"),
                  "fused_sgd_in_allreduce": self.fuse_sgd, "gradient_exchange_dtype": self.gradient_dtype,
                  "sm_partition": self.reserved_sms if algo == "nvls_excl" else 0,
                  "small_bucket_algo": "one-shot", "comm_blocks": self.blocks, "nccl_calls_on_grad_path": 0})
        return d

    def bucket_sizes(self) -> List[int]:
        from ..models.engine import SMALL_NUMEL
        return [10 * self.engine.F * (2 if self.gradient_dtype == "bf16" else 4), SMALL_NUMEL * 4]

    # ------------------------------------------------------------ per step --
    def before_forward(self) -> None:
        """C4: BatchNorm running stats follow rank 0 (they are not used by the train-mode
        forward; this keeps checkpoints identical on every rank, like DDP does)."""
        # Normally nothing to do: the broadcast of step k's buffers is folded into step k's small-bucket kernel
        # (small_grads_ready), which leaves exactly the state DDP's broadcast before forward k+1 would.  A forward
        # that was not preceded by such a step (the first one after a load, two forwards in a row) syncs here.
        if self.world > 1 and self.broadcast_buffers and not self._bn_synced:
            self.backend.broadcast_symmetric_inplace(self.engine.bn_flat, 0)
        self._bn_synced = False

    def fc_grad_ready(self, engine) -> None:
        """Called right after the fc backward kernel is enqueued (main stream)."""
        if self.world == 1 or self._debug_skip_fc:
            return
        main = torch.cuda.current_stream()
        ready = torch.cuda.Event()
        ready.record(main)
        if self._onepass_ready is not None:
            from .. import ops
            with torch.cuda.stream(self.comm_stream), trace.nvtx_range("b200dp.allreduce.fc_bucket(onepass)"):
                self.comm_stream.wait_event(ready)
                self.backend.all_reduce_tiles(engine.fcw_grad, engine.fcw, self._onepass_ready, rows=10,
                                              slab_floats=ops.fc_bwd_tile_slab_floats(8), target=0,
                                              scale=1.0 / self.world, lr=engine.lr, channel=1,
                                              blocks=max(1, self.reserved_sms or 8),
                                              use_nvls=None if self.algo in (None, "nvls", "nvls_excl") else False)
                self._fc_event = torch.cuda.Event()
                self._fc_event.record(self.comm_stream)
            return
        with torch.cuda.stream(self.comm_stream), trace.nvtx_range("b200dp.allreduce.fc_bucket"):
            self.comm_stream.wait_event(ready)
            if self.gradient_dtype == "bf16":
                self.backend.all_reduce_bf16_sgd(engine.fcw_grad, engine.fcw, engine.lr, 1.0 / self.world, channel=1,
                                                 blocks=self.blocks)
            else:
                self.backend.all_reduce(engine.fcw_grad, dist.ReduceOp.SUM, channel=1, scale=1.0 / self.world,
                                        sgd=(engine.fcw, engine.lr) if self.fuse_sgd else None, algo=self.algo,
                                        blocks=self.blocks)
            self._fc_event = torch.cuda.Event()
            self._fc_event.record(self.comm_stream)

    def fc_bwd_tiles(self, engine, dlogits, p2, dp2, bn2, bwd_partials) -> bool:
        """Tile pipeline: launch the producing fc backward on the main stream and the consuming all-reduce beside it
        on the side stream.  Returns False when the pipeline is not in use (caller runs the plain fc backward)."""
        if self.world == 1 or self.tile_ready is None or p2.shape[0] > 8 or self._debug_skip_fc:
            return False
        from .. import ops
        main = torch.cuda.current_stream()
        start = torch.cuda.Event()
        start.record(main)                                   # BEFORE the fc backward: the two kernels run concurrently
        self._tile_steps += 1
        # producer first: once its launch call has returned it cannot be held up behind the spinning consumer
        ops.fc_bwd_tiles(dlogits, p2, engine.fcw, engine.fcw_grad, dp2, self.tile_ready, self.tile_sweeps,
                         gscale=engine.grad_scale, bn2=bn2, bwd_partials=bwd_partials)
        with torch.cuda.stream(self.comm_stream), trace.nvtx_range("b200dp.allreduce.fc_bucket(tiles)"):
            self.comm_stream.wait_event(start)
            self.backend.all_reduce_tiles(engine.fcw_grad, engine.fcw, self.tile_ready, rows=10,
                                          slab_floats=self.tile_slab_floats,
                                          target=self._tile_steps * self.tile_blocks, scale=1.0 / self.world,
                                          lr=engine.lr, channel=1, blocks=max(1, self.reserved_sms or 8),
                                          use_nvls=None if self.algo in (None, "nvls", "nvls_excl") else False)
            self._fc_event = torch.cuda.Event()
            self._fc_event.record(self.comm_stream)
        return True

    def small_grads_ready(self, engine) -> bool:
        """One-shot all-reduce of the 53 KB bucket with the SGD update fused in; True = parameters updated."""
        if self.world == 1:
            return False
        fold = self.broadcast_buffers and os.environ.get("B200DP_FOLD_BN_BCAST", "1") == "1"
        self.backend.all_reduce_small_symmetric(engine.small_grad, 1.0 / self.world, sgd=(engine.small, engine.lr),
                                                bcast=(engine.bn_flat, 0) if fold else None)
        self._bn_synced = fold
        return True

    def wait_fc_grad(self, engine) -> bool:
        """Main stream waits for the bucket all-reduce; True when SGD was applied in it."""
        if self.world == 1:
            return False
        if self._fc_event is not None:
            torch.cuda.current_stream().wait_event(self._fc_event)
            self._fc_event = None
        self._steps += 1
        return self.fuse_sgd
