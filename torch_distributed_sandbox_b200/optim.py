"""Optimizers.  The reference uses exactly one: plain SGD, lr 1e-4 (mnist_onegpu.py:49)."""
from .models.module import SGD  # noqa: F401
