"""``FusedModule`` - an ``nn.Module`` face for the fused engine so that the reference's
training-loop idiom keeps working unchanged (mnist_onegpu.py:63-74):

    outputs = model(images); loss = criterion(outputs, labels)
    optimizer.zero_grad(); loss.backward(); optimizer.step()

``model(images)`` runs the engine forward and returns logits attached to autograd through
a custom Function whose backward launches the engine's fused backward; parameters are
``nn.Parameter`` views over the engine's buffers whose ``.grad`` point at the engine's
gradient buffers, so ``torch.optim.SGD`` works too (``optim.SGD`` here is the fused one).
``train_step(images, labels)`` is the fast path (loss fused with the fc tail, SGD fused
into the fc backward) that bench.py and the entrypoints use.
"""
from __future__ import annotations

from collections import OrderedDict

import torch
from torch import nn

from .engine import FusedConvNet


class _EngineFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, anchor, images, engine):
        ctx.engine = engine
        return engine.forward(images).clone()

    @staticmethod
    def backward(ctx, dlogits):
        eng = ctx.engine
        eng.backward_from_dlogits(dlogits.contiguous())
        eng._publish_grads()
        return torch.zeros((), device=dlogits.device), None, None


class FusedModule(nn.Module):
    def __init__(self, image_shape=(3000, 3000), batch_size: int = 5, device=None, lr: float = 1e-4, seed: int = 0,
                 **engine_kwargs):
        super().__init__()
        self.engine = FusedConvNet(image_shape, batch_size=batch_size, device=device, lr=lr, seed=seed,
                                   **engine_kwargs)
        eng = self.engine
        self._anchor = nn.Parameter(torch.zeros((), device=eng.device))       # ties the graph to autograd
        self._named = OrderedDict()
        for k, v in eng.params.items():
            self._named[k] = nn.Parameter(v, requires_grad=True)
        self._named["fc.weight"] = nn.Parameter(eng.fcw, requires_grad=True)   # NOTE: HWC-flatten order
        eng._publish_grads = self._publish_grads
        eng._on_rebind = self._on_rebind

    def _on_rebind(self):
        self._named["fc.weight"] = nn.Parameter(self.engine.fcw, requires_grad=True)

    def _publish_grads(self):
        eng = self.engine
        for k, p in self._named.items():
            p.grad = eng.fcw_grad if k == "fc.weight" else eng.grads[k]

    # nn.Module protocol ---------------------------------------------------------
    def named_parameters(self, prefix: str = "", recurse: bool = True, remove_duplicate: bool = True):
        for k, p in self._named.items():
            yield (prefix + "." if prefix else "") + k, p

    def parameters(self, recurse: bool = True):
        for _, p in self.named_parameters():
            yield p

    def state_dict(self, *args, destination=None, prefix: str = "", keep_vars: bool = False):
        sd = self.engine.state_dict(prefix=prefix)
        if destination is not None:
            destination.update(sd)
            return destination
        return sd

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        self.engine.load_state_dict(state_dict, strict=strict)

    def forward(self, images: torch.Tensor) -> torch.Tensor:
        if torch.is_grad_enabled():
            return _EngineFn.apply(self._anchor, images, self.engine)
        return self.engine.forward(images).clone()

    def train_step(self, images: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        return self.engine.train_step(images, labels)


class SGD:
    """Fused SGD over a FusedModule (``p -= lr * g``; the reference's
    ``torch.optim.SGD(model.parameters(), 1e-4)``, mnist_onegpu.py:49)."""

    def __init__(self, module, lr: float = 1e-4):
        while not isinstance(module, FusedModule):
            module = module.module                     # unwrap DistributedDataParallel
        self.engine = module.engine
        self.engine.lr = float(lr)
        self.engine._fc_fused_done = False

    def zero_grad(self, set_to_none: bool = True):
        pass                                           # gradient buffers are overwritten, never accumulated

    def step(self):
        self.engine.step()
