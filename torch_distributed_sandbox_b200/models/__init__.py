"""Model family of the sandbox: the reference CNN as an eager fp32 module (CPU oracle,
gloo plumbing runs) and as the fused sm_100a training engine."""
from .reference_net import ConvNet, reference_init, feature_size, STATE_KEYS  # noqa: F401


def __getattr__(name):
    if name == "FusedConvNet":
        from .engine import FusedConvNet
        return FusedConvNet
    if name in ("FusedModule", "SGD"):
        from . import module
        return getattr(module, name)
    raise AttributeError(name)
