"""FusedConvNet - the B200-native training engine for the sandbox CNN.

One object owns every parameter, buffer, activation and scratch tensor of the
model (nothing is allocated inside a step) and runs the whole training step -
resize, forward, loss, backward, optimizer - as ~22 hand-written sm_100a kernel
launches on one stream, instead of the reference's autograd graph of cuDNN /
ATen / cuBLAS calls (mnist_onegpu.py:63-74, mnist_distributed.py:85-97;
SURVEY.md 3.5 and 2.5 K1-K15).

Data flow (bf16 NHWC activations, fp32 parameters and statistics):

    u8 28x28 --resize--> x fp32 [B,H,W]
      --conv1 (FFMA, +BN1 stat partials)--> x1 [B,H,W,16]
      --BN1+ReLU+pool--> p1 [B,H/2,W/2,16]
      --conv2 (tcgen05, +BN2 stat partials)--> x2 [B,H/2,W/2,32]
      --BN2+ReLU+pool--> p2 [B,H/4,W/4,32] == fc input (HWC flatten)
      --fc (weight streaming) + cross-entropy--> loss, dlogits
    backward: fc_bwd (dW | fused SGD, dp2) -> pool/ReLU/BN2 bwd (2 passes, arg-max
      recomputed) -> conv2 wgrad + dgrad (tcgen05) -> BN1 stats -> fused
      pool/ReLU/BN1-bwd + conv1 wgrad -> SGD.

Gradient buffers are written IN PLACE over the activations they replace
(dp2 over p2, dx2 over x2, dp1 over p1), so resident memory is ~0.58 GB/image at
3000x3000 against ~2.4 GB/image retained by the reference: the
"OOM workaround" headroom of README.md:9-15 on a 180 GB part.

In data-parallel mode a ``comm`` object (parallel/ddp.py::EngineGradSync) is
attached: fc_bwd then writes the 720 MB fc.weight gradient, pre-divided by the
world size, straight into the symmetric-memory bucket and the all-reduce kernel
runs on a side stream while the conv backward continues (K10 -> K16).
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Optional, Sequence

import torch
import torch.nn.functional as F

from ..utils import trace

from .. import ops
from ..data.resize import pil_bilinear_table
from . import reference_net as rn

# (state_dict key, shape) of the parameters kept in the "small" flat buffer, in order.
SMALL_LAYOUT = (
    ("layer1.0.weight", (16, 1, 5, 5)), ("layer1.0.bias", (16,)),
    ("layer1.1.weight", (16,)), ("layer1.1.bias", (16,)),
    ("layer2.0.weight", (32, 16, 5, 5)), ("layer2.0.bias", (32,)),
    ("layer2.1.weight", (32,)), ("layer2.1.bias", (32,)),
    ("fc.bias", (10,)),
)


def _numel(shape) -> int:
    n = 1
    for s in shape:
        n *= s
    return n


SMALL_NUMEL = sum(_numel(s) for _, s in SMALL_LAYOUT)          # 13,354
SMALL_ALLOC = (SMALL_NUMEL + 3) // 4 * 4


class FusedConvNet:
    """Fused training engine.  ``train_step(images, labels)`` returns the loss (0-dim CUDA tensor)."""

    def __init__(self, image_shape: Sequence[int] = (3000, 3000), batch_size: int = 5,
                 device: Optional[torch.device] = None, lr: float = 1e-4, seed: int = 0,
                 init: str = "reference", conv_backend: str = "tc", fuse_fc_sgd: bool = True,
                 precision: Optional[str] = None,
                 fc_weight_storage: Optional[torch.Tensor] = None,
                 fc_grad_storage: Optional[torch.Tensor] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("FusedConvNet needs a CUDA device (sm_100a); use models.ConvNet on CPU")
        H, W = (int(v) for v in image_shape)
        if H % 4 or W % 4:
            raise ValueError("image height/width must be multiples of 4 (two 2x2 pools)")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.H, self.W, self.B = H, W, int(batch_size)
        self.image_shape = (H, W)
        self.lr = float(lr)
        self.conv_backend = conv_backend
        # "tf32": fp32 activation / gradient storage, kind::tf32 tensor-core math in conv2, hi/lo-split (fp32-level)
        #         products in conv1, fp32 everywhere else - the reference's own numerics (fp32 tensors, cuDNN with
        #         allow_tf32, fp32 cuBLAS fc; mnist_distributed.py:28-38 has no autocast).
        # "bf16": bf16 activation / gradient storage and tensor-core operands, fp32 accumulation: the fast mode.
        self.precision = (precision or os.environ.get("B200DP_PRECISION", "bf16")).lower()
        if self.precision in ("fp32", "float32"):
            self.precision = "tf32"
        if self.precision not in ("bf16", "tf32"):
            raise ValueError(f"precision must be 'bf16' or 'tf32', got {precision!r}")
        self.act_dtype = torch.float32 if self.precision == "tf32" else torch.bfloat16
        self.conv1_backend = os.environ.get("B200DP_CONV1", "tc")      # "tc" (tcgen05) | "ffma"
        # BN-backward reductions fused into fc_bwd / the conv2-dgrad epilogue (0: separate passes over x2 / x1)
        self.fuse_bwd_stats = os.environ.get("B200DP_FUSE_BWD_STATS", "1") == "1"
        self.fuse_fc_sgd = bool(fuse_fc_sgd)
        self.conv1_two_pass = os.environ.get("B200DP_CONV1_TWO_PASS", "1") == "1"    # tf32 mode only
        # tf32 mode: never store the conv1 output (2.88 GB at batch 5).  The second forward pass writes the pooled tensor
        # and one byte per pooled element (arg-max | ReLU), the weight-gradient kernel routes dp1 with those bytes and
        # gets the dense BatchNorm term from the input's Gram matrix.
        self.conv1_no_x1 = (os.environ.get("B200DP_CONV1_NO_X1", "1") == "1" and self.precision == "tf32" and
                            self.conv1_two_pass and self.fuse_bwd_stats and conv_backend == "tc" and
                            H % 8 == 0 and W % 4 == 0)
        # Layer-1 forward kernels of the two-pass path.  "vt" (default): bf16 hi/lo vertical-Toeplitz kernel, fp32-level
        # products - what the reference effectively computes at 3000x3000, where cuDNN picks fp32 algorithms.  "hz": the
        # kind::tf32 kernel that takes the image rows themselves as its MMA operand (csrc/conv1_hz.cu), 0.29 ms faster per
        # step; a 2^-11 operand rounding in THIS conv is amplified by the BatchNorm-backward cancellation to tens of percent
        # in single-step conv-weight gradients at >= 1000x1000 (profiles/precision_error_vs_cudnn_tf32.txt), loss curves
        # are unaffected (profiles/convergence_3000_lr1e-6_hz.txt) - so it is the opt-in.
        self.conv1_row_operand = self.conv1_no_x1 and os.environ.get("B200DP_CONV1_FWD", "vt") == "hz"
        self.F = rn.feature_size((H, W))
        self.comm = None           # set by parallel.ddp for data-parallel runs
        self._graph_enabled = os.environ.get("B200DP_CUDA_GRAPH", "1") == "1"
        self._graphs = {}          # (batch, src h, src w, lr, grad scale) -> (CUDAGraph | None, launches | eager count, key)
        self._b = int(batch_size)  # size of the batch currently in the buffers (<= B: short last batch)
        self._fc_fused_done = False
        self._fc_pending = False
        self._publish_grads = lambda: None
        self._on_rebind = lambda: None
        self.grad_scale = 1.0      # 1/world_size under DDP (fused into the producing kernels)
        dev, B = self.device, self.B
        f32 = dict(device=dev, dtype=torch.float32)
        bf = dict(device=dev, dtype=self.act_dtype)

        with torch.cuda.device(dev):
            # ---- parameters / buffers --------------------------------------
            self.small = torch.zeros(SMALL_ALLOC, **f32)
            self.small_grad = torch.zeros(SMALL_ALLOC, **f32)
            self.params: "OrderedDict[str, torch.Tensor]" = OrderedDict()
            self.grads: "OrderedDict[str, torch.Tensor]" = OrderedDict()
            off = 0
            for name, shape in SMALL_LAYOUT:
                n = _numel(shape)
                self.params[name] = self.small[off:off + n].view(shape)
                self.grads[name] = self.small_grad[off:off + n].view(shape)
                off += n
            self.fcw = fc_weight_storage if fc_weight_storage is not None else torch.empty(10, self.F, **f32)
            assert self.fcw.shape == (10, self.F) and self.fcw.dtype == torch.float32
            self.fcw_grad = fc_grad_storage      # allocated lazily unless fused SGD covers it
            # BatchNorm buffers live in ONE 512-byte blob so that DDP's per-forward buffer
            # broadcast (C4) is a single tiny kernel: [rm1 16 | rv1 16 | rm2 32 | rv2 32 | nbt1 | nbt2]
            self.bn_flat = torch.zeros(128, **f32)
            self.bn_flat[16:32] = 1.0
            self.bn_flat[64:96] = 1.0
            self._bind_bn_views()
            # ---- activations (gradients alias them) -------------------------
            self.x = torch.empty(B, H, W, **f32)
            self.x1 = None if self.conv1_no_x1 else torch.empty(B, H, W, 16, **bf)
            self.mask1 = torch.empty(B, H // 2, W // 2, 16, device=dev, dtype=torch.uint8) if self.conv1_no_x1 else None
            self.p1 = torch.empty(B, H // 2, W // 2, 16, **bf)
            self.x2 = torch.empty(B, H // 2, W // 2, 32, **bf)
            self.p2 = torch.empty(B, H // 4, W // 4, 32, **bf)
            self.dp2, self.dx2, self.dp1 = self.p2, self.x2, self.p1     # in-place gradients
            # ---- scratch ------------------------------------------------------
            self.g_c1, self.g_wg1 = ops.conv1_fwd_grid(), ops.conv1_wgrad_grid()
            self.g_tc, self.g_pbb, self.g_fc = ops.conv_tc_max_grid(), ops.pool_bn_bwd_grid(), ops.fc_grid()
            self.part_c1 = torch.empty(self.g_c1, 32, **f32)
            self.part_c1b = torch.empty(self.g_c1, 32, **f32)       # scratch of the second conv1 pass (unused sums)
            self.part_c2 = torch.empty(self.g_tc, 64, **f32)
            self.part_bwd = torch.empty(self.g_pbb, 64, **f32)
            self.part_wg1 = torch.empty(self.g_wg1, ops.conv1_wgrad_nox1_partial() if self.conv1_no_x1 else 400, **f32)
            self.part_wg2 = torch.empty(self.g_tc, 32 * 16 * 25, **f32)
            self.part_fc = torch.empty(self.g_fc, 80, **f32)
            self.part_fcb = torch.empty(self.g_fc, 64, **f32)
            self.bn1 = torch.zeros(4, 16, **f32)
            self.bn2 = torch.zeros(4, 32, **f32)
            self.red1 = torch.zeros(2, 16, **f32)
            self.red2 = torch.zeros(2, 32, **f32)
            self.logits = torch.zeros(B, 10, **f32)
            self.dlogits = torch.zeros(B, 10, **f32)
            self.loss = torch.zeros((), **f32)
            self._tabs = {}
            self._u8 = None
            self._labels = torch.zeros(B, device=dev, dtype=torch.int64)

        if init == "reference":
            self.load_state_dict(rn.reference_init((H, W), seed=seed).state_dict())
        elif init == "random":
            g = torch.Generator(device="cpu").manual_seed(seed)
            sd = OrderedDict()
            for name, shape in SMALL_LAYOUT:
                fan = _numel(shape[1:]) if len(shape) > 1 else 1
                if name.endswith("1.weight"):      # BN gamma
                    sd[name] = 1.0 + 0.1 * torch.randn(shape, generator=g)
                elif len(shape) == 1:
                    sd[name] = 0.1 * torch.randn(shape, generator=g)
                else:
                    sd[name] = torch.randn(shape, generator=g) / fan ** 0.5
            sd["fc.weight"] = torch.randn(10, self.F, generator=g) / self.F ** 0.5
            self.load_state_dict(sd, strict=False)
        elif init != "empty":
            raise ValueError(f"unknown init {init!r}")

    def _bind_bn_views(self) -> None:
        f = self.bn_flat
        self.bn_buf = {
            1: dict(mean=f[0:16], var=f[16:32], nbt=f[96:98].view(torch.int64)[0]),
            2: dict(mean=f[32:64], var=f[64:96], nbt=f[98:100].view(torch.int64)[0]),
        }

    def rebind_storage(self, fcw: Optional[torch.Tensor] = None, fcw_grad: Optional[torch.Tensor] = None,
                       bn_flat: Optional[torch.Tensor] = None, small_grad: Optional[torch.Tensor] = None) -> None:
        """Move fc.weight / its gradient / the BN buffer blob into caller-provided storage
        (symmetric memory under DDP); current values are preserved."""
        self._graphs.clear()                # captured steps address the old storage
        if fcw is not None:
            fcw = fcw.view(10, self.F)
            fcw.copy_(self.fcw)
            self.fcw = fcw
        if fcw_grad is not None:
            self.fcw_grad = fcw_grad.view(10, self.F)
        if small_grad is not None:
            small_grad.zero_()      # the conv-bias slots are never written again (analytically zero gradients)
            self.small_grad = small_grad
            off = 0
            for name, shape in SMALL_LAYOUT:
                n = _numel(shape)
                self.grads[name] = self.small_grad[off:off + n].view(shape)
                off += n
        if bn_flat is not None:
            bn_flat.copy_(self.bn_flat)
            self.bn_flat = bn_flat
            self._bind_bn_views()
        self._on_rebind()

    # ------------------------------------------------------------ checkpoint --
    def state_dict(self, prefix: str = "") -> "OrderedDict[str, torch.Tensor]":
        """torch ``state_dict`` with the reference's 16 keys; ``fc.weight`` is
        converted from the internal HWC-flatten order back to the reference's CHW order."""
        out = OrderedDict()
        self._sync_fc()
        with torch.cuda.device(self.device):
            fcw_chw = torch.empty_like(self.fcw)
            ops.fcw_permute(self.fcw, fcw_chw, 32, self.H // 4, self.W // 4, to_hwc=False)
        for k in rn.STATE_KEYS:
            if k == "fc.weight":
                v = fcw_chw
            elif k in self.params:
                v = self.params[k].clone()
            else:
                layer = 1 if k.startswith("layer1") else 2
                field = {"running_mean": "mean", "running_var": "var", "num_batches_tracked": "nbt"}[k.split(".")[-1]]
                v = self.bn_buf[layer][field].clone()
            out[prefix + k] = v
        return out

    def load_state_dict(self, sd, strict: bool = True) -> None:
        """Accepts a reference/torch ``state_dict`` (optionally with DDP's ``module.`` prefix)."""
        sd = {(k[7:] if k.startswith("module.") else k): v for k, v in sd.items()}
        missing = [k for k in rn.STATE_KEYS if k not in sd]
        if strict and missing:
            raise KeyError(f"state_dict is missing {missing}")
        with torch.cuda.device(self.device), torch.no_grad():
            for k, v in sd.items():
                if k == "fc.weight":
                    if tuple(v.shape) != (10, self.F):
                        raise ValueError(f"fc.weight shape {tuple(v.shape)} != {(10, self.F)}")
                    src = v.to(self.device, torch.float32).contiguous()
                    ops.fcw_permute(src, self.fcw, 32, self.H // 4, self.W // 4, to_hwc=True)
                    torch.cuda.current_stream().synchronize()
                elif k in self.params:
                    self.params[k].copy_(v.to(self.device, torch.float32))
                elif k in rn.STATE_KEYS:
                    layer = 1 if k.startswith("layer1") else 2
                    field = {"running_mean": "mean", "running_var": "var", "num_batches_tracked": "nbt"}[k.split(".")[-1]]
                    self.bn_buf[layer][field].copy_(v.to(self.device))
                elif strict:
                    raise KeyError(f"unexpected key {k}")

    def named_gradients(self) -> "OrderedDict[str, torch.Tensor]":
        """Gradients of the last backward in the reference's layouts (tests / generic DDP)."""
        out = OrderedDict((k, v.clone()) for k, v in self.grads.items())
        if self.fcw_grad is not None:
            src = self.fcw_grad.float() if self.fcw_grad.dtype != torch.float32 else self.fcw_grad   # bf16 DDP bucket
            g = torch.empty_like(src)
            ops.fcw_permute(src, g, 32, self.H // 4, self.W // 4, to_hwc=False)
            out["fc.weight"] = g
        return out

    # ------------------------------------------------------------------ input --
    def set_input(self, images: torch.Tensor) -> None:
        """uint8 [B,h,w] raw digits (resized on the GPU, K2) or float [B,1,H,W] / [B,H,W]."""
        if images.dtype == torch.uint8:
            if images.dim() == 4:
                images = images[:, 0]
            b = int(images.shape[0])
            assert 1 <= b <= self.B, f"batch {b} exceeds the engine's batch capacity {self.B}"
            self._b = b
            sh, sw = int(images.shape[-2]), int(images.shape[-1])
            key = (sh, sw)
            if key not in self._tabs:
                self._tabs[key] = (torch.from_numpy(pil_bilinear_table(sw, self.W)).to(self.device),
                                   torch.from_numpy(pil_bilinear_table(sh, self.H)).to(self.device))
            if self._u8 is None or self._u8.shape[1:] != images.shape[1:]:
                self._u8 = torch.empty((self.B,) + tuple(images.shape[1:]), device=self.device, dtype=torch.uint8)
                self._graphs.clear()        # captured steps read the old staging buffer
            self._u8[:b].copy_(images, non_blocking=True)
            xt, yt = self._tabs[key]
            ops.resize_pil_bilinear(self._u8[:b], xt, yt, self.x[:b])
        else:
            if images.dim() == 4:
                images = images[:, 0]
            b = int(images.shape[0])
            assert 1 <= b <= self.B and tuple(images.shape[1:]) == (self.H, self.W), \
                f"bad input shape {tuple(images.shape)} (capacity {self.B} x {self.H} x {self.W})"
            self._b = b
            self.x[:b].copy_(images, non_blocking=True)

    # ---------------------------------------------------------------- forward --
    def forward(self, images: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Train-mode forward (batch statistics, running stats updated). Returns logits [B,10]."""
        if images is not None:
            self.set_input(images)
        # B = size of THIS batch: a short last batch (the reference's loaders are drop_last=False,
        # mnist_onegpu.py:55-59) runs on leading-dimension views of the same buffers
        P, B, H, W = self.params, self._b, self.H, self.W
        x, p1, x2, p2t = self.x[:B], self.p1[:B], self.x2[:B], self.p2[:B]
        x1 = self.x1[:B] if self.x1 is not None else None
        two_pass = self.precision == "tf32" and self.conv1_two_pass
        if two_pass:
            # fp32 storage: conv1 is cheap on the tensor cores (0.24 ms without its store), its 2.88 GB output is not -
            # a statistics-only pass, then conv + store + BN/ReLU/pool in one epilogue instead of a pass over x1
            n1 = ops.conv1_stats_tc(x, P["layer1.0.weight"], P["layer1.0.bias"], self.part_c1,
                                    row_operand=self.conv1_row_operand)
        elif self.conv1_backend == "tc" or self.precision == "tf32":
            n1 = ops.conv1_fwd_tc(x, P["layer1.0.weight"], P["layer1.0.bias"], x1, self.part_c1)
        else:
            ops.conv1_fwd(x, P["layer1.0.weight"], P["layer1.0.bias"], x1, self.part_c1)
            n1 = self.g_c1
        ops.bn_finalize(self.part_c1, n1, 16, B * H * W, P["layer1.1.weight"], P["layer1.1.bias"],
                        self.bn_buf[1]["mean"], self.bn_buf[1]["var"], self.bn_buf[1]["nbt"], self.bn1)
        if two_pass:
            ops.conv1_fwd_pool_tc(x, P["layer1.0.weight"], P["layer1.0.bias"], self.bn1, x1, p1, self.part_c1b,
                                  amask_u8=self.mask1[:B] if self.conv1_no_x1 else None,
                                  row_operand=self.conv1_row_operand)
        else:
            ops.bn_relu_pool(x1, self.bn1, p1)
        if self.conv_backend == "tc":
            nparts = ops.conv2_fwd_tc(p1, P["layer2.0.weight"], P["layer2.0.bias"], x2, self.part_c2)
        else:
            nparts = self._conv2_fwd_torch()
        ops.bn_finalize(self.part_c2, nparts, 32, B * (H // 2) * (W // 2), P["layer2.1.weight"],
                        P["layer2.1.bias"], self.bn_buf[2]["mean"], self.bn_buf[2]["var"], self.bn_buf[2]["nbt"],
                        self.bn2)
        ops.bn_relu_pool(x2, self.bn2, p2t)
        p2 = p2t.view(B, self.F)
        self._sync_fc()
        for b0 in range(0, B, 8):
            b1 = min(B, b0 + 8)
            ops.fc_fwd(p2[b0:b1], self.fcw, P["fc.bias"], self.part_fc, self.logits[b0:b1])
        return self.logits[:B]

    # --------------------------------------------------------------- backward --
    def loss_backward(self, labels: torch.Tensor) -> torch.Tensor:
        """Cross-entropy (mean) + full backward.  Leaves gradients in ``self.grads`` /
        ``self.fcw_grad``; with ``fuse_fc_sgd`` and no comm, fc.weight is updated in place here."""
        B = self._b
        assert labels.shape[0] == B, f"{labels.shape[0]} labels for a batch of {B}"
        self._labels[:B].copy_(labels, non_blocking=True)
        ops.cross_entropy(self.logits[:B], self._labels[:B], self.loss, self.dlogits[:B], self.grads["fc.bias"],
                          self.grad_scale)
        self._backward(allow_fused_sgd=True)
        return self.loss

    def backward_from_dlogits(self, dlogits: torch.Tensor) -> None:
        """Backward for an externally computed loss (autograd front-end, models/module.py)."""
        B = self._b
        self.dlogits[:B].copy_(dlogits)
        self.grads["fc.bias"].copy_(self.dlogits[:B].sum(0) * self.grad_scale)
        self._backward(allow_fused_sgd=False)

    def _backward(self, allow_fused_sgd: bool) -> None:
        P, G, B, H, W = self.params, self.grads, self._b, self.H, self.W
        x, p1, x2 = self.x[:B], self.p1[:B], self.x2[:B]
        x1 = self.x1[:B] if self.x1 is not None else None
        dx2, dp1 = x2, p1                  # in-place gradients (dp2 over p2, dx2 over x2, dp1 over p1)
        gs = self.grad_scale
        fused = allow_fused_sgd and self.fuse_fc_sgd and self.comm is None
        if not fused and self.fcw_grad is None:
            self.fcw_grad = torch.empty_like(self.fcw)
        p2 = self.p2[:B].view(B, self.F)
        dp2 = p2
        dp2t = self.p2[:B]
        fuse_stats = self.fuse_bwd_stats and self.conv_backend == "tc"
        tiled = (self.comm is not None and hasattr(self.comm, "fc_bwd_tiles") and
                 self.comm.fc_bwd_tiles(self, self.dlogits[:B], p2, dp2, self.bn2 if fuse_stats else None,
                                        self.part_fcb if fuse_stats else None))
        if not tiled:
            # one launch for any batch: <= 8 rows stay in registers, larger batches loop inside the single pass over W
            ops.fc_bwd(self.dlogits[:B], p2, self.fcw, None if fused else self.fcw_grad, dp2,
                       lr=self.lr if fused else 0.0, accumulate=False, gscale=gs,
                       bn2=self.bn2 if fuse_stats else None, bwd_partials=self.part_fcb if fuse_stats else None)
        self._fc_fused_done = fused
        # async all-reduce of the 720 MB bucket.  By default it starts after the next kernel instead of here:
        # pool_bn_bwd is HBM-bound and lost 0.16 ms to the all-reduce's 1.4 GB of HBM traffic, while the
        # conv kernels after it are bound by the shared-memory pipe / tensor pipe and have DRAM headroom.
        defer_ar = self.comm is not None and getattr(self.comm, "start_after_pool", False) and not tiled
        if self.comm is not None and not defer_ar and not tiled:
            self.comm.fc_grad_ready(self)
        # ---- layer 2 (statistics pass fused into fc_bwd above unless disabled)
        H2, W2 = H // 2, W // 2
        if fuse_stats:
            ops.bn_bwd_finalize(self.part_fcb, self.g_fc, 32, B * H2 * W2, self.red2, G["layer2.1.weight"],
                                G["layer2.1.bias"], gs)
        else:
            ops.pool_bn_bwd(x2, dp2t, self.bn2, self.red2, self.part_bwd, None, mode=0)
            ops.bn_bwd_finalize(self.part_bwd, self.g_pbb, 32, B * H2 * W2, self.red2, G["layer2.1.weight"],
                                G["layer2.1.bias"], gs)
        ops.pool_bn_bwd(x2, dp2t, self.bn2, self.red2, None, dx2, mode=1)
        if defer_ar:
            self.comm.fc_grad_ready(self)
        if self.conv_backend == "tc":
            ops.conv2_wgrad_tc(p1, dx2, self.part_wg2, G["layer2.0.weight"], gs)
            if fuse_stats:   # dgrad epilogue also reduces BN1's backward sums (reads p1 before dp1 overwrites it)
                n1 = ops.conv2_dgrad_stats_tc(dx2, P["layer2.0.weight"], p1, self.bn1, dp1, self.part_bwd)
            else:
                ops.conv2_dgrad_tc(dx2, P["layer2.0.weight"], dp1)
        else:
            self._conv2_bwd_torch(gs)
            G["layer2.0.bias"].zero_()
        # a bias feeding BatchNorm has an analytically zero gradient: its slot of the gradient buffer is zeroed once, when
        # the buffer is created, and no native kernel writes it (two fill launches per step less)
        # ---- layer 1
        if fuse_stats:
            ops.bn_bwd_finalize(self.part_bwd, n1, 16, B * H * W, self.red1, G["layer1.1.weight"],
                                G["layer1.1.bias"], gs)
        else:
            ops.pool_bn_bwd(x1, dp1, self.bn1, self.red1, self.part_bwd, None, mode=0)
            ops.bn_bwd_finalize(self.part_bwd, self.g_pbb, 16, B * H * W, self.red1, G["layer1.1.weight"],
                                G["layer1.1.bias"], gs)
        if self.conv1_no_x1:
            ops.conv1_wgrad_nox1(x, dp1, self.mask1[:B], P["layer1.0.weight"], P["layer1.0.bias"], self.bn1, self.red1,
                                 self.part_wg1, G["layer1.0.weight"], gs)
        else:
            wg1 = ops.conv1_wgrad_tc if self.conv1_backend == "tc" else ops.conv1_wgrad
            wg1(x, x1, dp1, self.bn1, self.red1, self.part_wg1, G["layer1.0.weight"], gs)

    def step(self) -> None:
        """SGD (lr, no momentum / weight decay: mnist_onegpu.py:49) on every parameter.  Under DDP the
        small bucket's update is fused into its one-shot all-reduce, and the fc.weight update (fused into
        the overlapped bucket all-reduce) is only WAITED FOR right before the next fc forward
        (``_sync_fc``): nothing earlier in the next step reads fc.weight, so the all-reduce also hides
        under the next forward's conv layers."""
        small_done = self.comm.small_grads_ready(self) if self.comm is not None else False
        if not small_done:
            ops.sgd(self.small, self.small_grad, self.lr)
        if self.comm is not None:
            self._fc_pending = True
        elif not self._fc_fused_done:
            self._sgd_fcw()

    def _sgd_fcw(self) -> None:
        if self.fcw_grad.dtype == torch.float32:
            ops.sgd(self.fcw, self.fcw_grad, self.lr)
        else:                                              # bf16 gradient bucket without the fused all-reduce epilogue
            self.fcw.add_(self.fcw_grad.float(), alpha=-self.lr)

    def _sync_fc(self) -> None:
        """Make fc.weight current (DDP: wait for the bucket all-reduce; apply SGD if it was not fused)."""
        if self._fc_pending:
            self._fc_pending = False
            if not self.comm.wait_fc_grad(self):
                self._sgd_fcw()

    def train_step(self, images: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        if not trace.nvtx_enabled():
            if self._graph_enabled and self.comm is None and images.dtype == torch.uint8:
                return self._train_step_graphed(images, labels)
            self.forward(images)
            self.loss_backward(labels)
            self.step()
            return self.loss
        with trace.nvtx_range("b200dp.forward"):
            self.forward(images)
        with trace.nvtx_range("b200dp.loss+backward"):
            self.loss_backward(labels)
        with trace.nvtx_range("b200dp.optimizer"):
            self.step()
        return self.loss

    # ------------------------------------------------------------ CUDA graph --
    def _step_body(self, b: int, key) -> None:
        """The whole step on static buffers: resize of the staged digits, forward, loss, backward, SGD."""
        xt, yt = self._tabs[key]
        ops.resize_pil_bilinear(self._u8[:b], xt, yt, self.x[:b])
        self.forward()
        ops.cross_entropy(self.logits[:b], self._labels[:b], self.loss, self.dlogits[:b], self.grads["fc.bias"],
                          self.grad_scale)
        self._backward(allow_fused_sgd=True)
        self.step()

    def _train_step_graphed(self, images: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """Single-GPU step as ONE graph launch (the ~20 kernels of a step on static buffers; the step's inputs are
        copied into the staging tensors first).  A step that follows a host synchronisation - every end-to-end step
        does: the loss is read back - otherwise starts with the GPU idle while Python issues the launches one by one.
        Two eager steps per (batch, source size, lr) come first; a failed capture disables the path for good.
        (Under DDP the collectives carry a per-step epoch argument, so those steps stay eager.)"""
        if images.dim() == 4:
            images = images[:, 0]
        b = int(images.shape[0])
        gkey = (b, int(images.shape[-2]), int(images.shape[-1]), float(self.lr), float(self.grad_scale))
        ent = self._graphs.get(gkey)
        if ent is None or ent[0] is None:
            n = 0 if ent is None else ent[1]
            if n < 2 or not (1 <= b <= self.B):
                self._graphs[gkey] = (None, n + 1)
                self.forward(images)
                self.loss_backward(labels)
                self.step()
                return self.loss
            key = (int(images.shape[-2]), int(images.shape[-1]))
            try:
                self._u8[:b].copy_(images, non_blocking=True)
                self._labels[:b].copy_(labels, non_blocking=True)
                self._b = b
                torch.cuda.current_stream().synchronize()
                graph = torch.cuda.CUDAGraph()
                l0 = ops.LAUNCHES
                with torch.cuda.graph(graph):
                    self._step_body(b, key)
                ent = (graph, ops.LAUNCHES - l0, key)
                ops._count(-(ops.LAUNCHES - l0))          # capture launched nothing
                self._graphs[gkey] = ent
            except Exception as e:                        # noqa: BLE001 - any capture failure: stay eager
                self._graph_enabled = False
                import warnings
                warnings.warn(f"b200dp: CUDA-graph capture of the training step failed ({e}); running eagerly")
                self.forward(images)
                self.loss_backward(labels)
                self.step()
                return self.loss
        graph, nlaunch, _ = ent
        assert labels.shape[0] == b, f"{labels.shape[0]} labels for a batch of {b}"
        self._b = b
        self._u8[:b].copy_(images, non_blocking=True)
        self._labels[:b].copy_(labels, non_blocking=True)
        graph.replay()
        ops._count(nlaunch)                               # the kernels of the replayed step
        return self.loss

    __call__ = forward

    # ------------------------------------------------ torch fallbacks (debug) --
    def _conv2_fwd_torch(self) -> int:
        P = self.params
        B = self._b
        x = self.p1[:B].permute(0, 3, 1, 2).float()
        w = P["layer2.0.weight"].to(torch.bfloat16).float()
        y = F.conv2d(x, w, P["layer2.0.bias"], padding=2)
        self.x2[:B].copy_(y.permute(0, 2, 3, 1))
        r = self.x2[:B].float()
        self.part_c2.zero_()
        self.part_c2[0, :32] = r.sum(dim=(0, 1, 2))
        self.part_c2[0, 32:] = (r * r).sum(dim=(0, 1, 2))
        return 1

    def _conv2_bwd_torch(self, gs: float) -> None:
        P, G = self.params, self.grads
        B = self._b
        dy = self.dx2[:B].permute(0, 3, 1, 2).float()
        x = self.p1[:B].permute(0, 3, 1, 2).float()
        w = P["layer2.0.weight"].to(torch.bfloat16).float()
        G["layer2.0.weight"].copy_(torch.nn.grad.conv2d_weight(x, w.shape, dy, padding=2) * gs)
        dxin = torch.nn.grad.conv2d_input(x.shape, w, dy, padding=2)
        self.dp1[:B].copy_(dxin.permute(0, 2, 3, 1))

    def activation_bytes(self) -> int:
        ts = (self.x, self.x1, self.mask1, self.p1, self.x2, self.p2)
        return sum(t.numel() * t.element_size() for t in ts if t is not None)
