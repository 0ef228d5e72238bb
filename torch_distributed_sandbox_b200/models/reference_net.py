"""Plain-PyTorch definition of the sandbox CNN, used as (a) the fp32 CPU oracle for
every fused kernel, (b) the model for the CPU/gloo plumbing configuration and
(c) the ``state_dict`` schema the fused engine imports/exports.

Architecture parity with the reference ``ConvNet`` (mnist_onegpu.py:11-31,
mnist_distributed.py:25-45): two ``conv5x5(pad 2) -> BatchNorm -> ReLU ->
maxpool2`` stages (1->16->32 channels) and a linear head whose input size
follows the image size.  Parameter/buffer names are the same 16 ``state_dict``
keys listed in SURVEY.md 2.4, so checkpoints move in both directions.

Unlike the reference, the head's fan-in is computed from ``image_shape`` instead
of being discovered by a 5.7 s CPU forward on zeros (mnist_onegpu.py:39);
``reference_init`` reproduces that forward's two side effects exactly: the RNG
order of the initialisers and the BatchNorm running-stat mutation.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Sequence

import torch
from torch import nn

CHANNELS = (1, 16, 32)
NUM_CLASSES = 10
STATE_KEYS = (
    "layer1.0.weight", "layer1.0.bias", "layer1.1.weight", "layer1.1.bias",
    "layer1.1.running_mean", "layer1.1.running_var", "layer1.1.num_batches_tracked",
    "layer2.0.weight", "layer2.0.bias", "layer2.1.weight", "layer2.1.bias",
    "layer2.1.running_mean", "layer2.1.running_var", "layer2.1.num_batches_tracked",
    "fc.weight", "fc.bias",
)


def feature_size(image_shape: Sequence[int]) -> int:
    h, w = image_shape
    return CHANNELS[2] * (h // 4) * (w // 4)


def _stage(cin: int, cout: int) -> nn.Sequential:
    return nn.Sequential(
        nn.Conv2d(cin, cout, kernel_size=5, stride=1, padding=2),
        nn.BatchNorm2d(cout),
        nn.ReLU(),
        nn.MaxPool2d(kernel_size=2, stride=2),
    )


class ConvNet(nn.Module):
    """Eager fp32 model with the reference's layer names (layer1 / layer2 / fc)."""

    def __init__(self, image_shape: Sequence[int] = (3000, 3000), num_classes: int = NUM_CLASSES):
        super().__init__()
        self.image_shape = tuple(int(v) for v in image_shape)
        self.layer1 = _stage(CHANNELS[0], CHANNELS[1])
        self.layer2 = _stage(CHANNELS[1], CHANNELS[2])
        self.fc = nn.Linear(feature_size(self.image_shape), num_classes)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        feats = self.layer2(self.layer1(x))
        return self.fc(feats.flatten(1))


def apply_lazy_init_side_effect(model: ConvNet) -> None:
    """Reproduce what the reference's train-mode forward on an all-zero image does to the
    BatchNorm buffers before training starts (mnist_onegpu.py:39).

    With zero input, conv1's output is its bias at every pixel, so BN1 sees mean = bias and
    variance 0: ``running_mean = 0.1 * bias``, ``running_var = 0.9``; BN1 then outputs beta = 0
    everywhere, so layer 2 sees zeros again and the same holds for BN2.  Both counters go to 1.
    """
    with torch.no_grad():
        for stage in (model.layer1, model.layer2):
            conv, bn = stage[0], stage[1]
            m = bn.momentum
            # torch computes (1 - m) * running + m * batch in fp32
            bn.running_mean.mul_(1 - m).add_(conv.bias.detach() * m)
            bn.running_var.mul_(1 - m)          # + m * 0
            bn.num_batches_tracked.add_(1)


def reference_init(image_shape: Sequence[int] = (3000, 3000), seed: int = 0,
                   lazy_side_effect: bool = True) -> ConvNet:
    """Model initialised exactly like the reference under ``torch.manual_seed(seed)``:
    conv1, conv2 (kaiming-uniform weight then bias), then the linear head - the order in which
    the reference consumes the global RNG, its LazyLinear being materialised last."""
    torch.manual_seed(seed)
    model = ConvNet(image_shape)
    if lazy_side_effect:
        apply_lazy_init_side_effect(model)
    return model


def empty_state_dict(image_shape: Sequence[int]) -> "OrderedDict[str, torch.Tensor]":
    """Shapes/dtypes of the 16 checkpoint entries (meta tensors; no memory)."""
    f = feature_size(image_shape)
    shapes = {
        "layer1.0.weight": (16, 1, 5, 5), "layer1.0.bias": (16,), "layer1.1.weight": (16,), "layer1.1.bias": (16,),
        "layer1.1.running_mean": (16,), "layer1.1.running_var": (16,), "layer1.1.num_batches_tracked": (),
        "layer2.0.weight": (32, 16, 5, 5), "layer2.0.bias": (32,), "layer2.1.weight": (32,), "layer2.1.bias": (32,),
        "layer2.1.running_mean": (32,), "layer2.1.running_var": (32,), "layer2.1.num_batches_tracked": (),
        "fc.weight": (NUM_CLASSES, f), "fc.bias": (NUM_CLASSES,),
    }
    out = OrderedDict()
    for k in STATE_KEYS:
        dt = torch.int64 if k.endswith("num_batches_tracked") else torch.float32
        out[k] = torch.empty(shapes[k], dtype=dt, device="meta")
    return out
