"""Dataset, loader and synthetic data (L5 / D9-D10 in SURVEY.md).

The reference reads MNIST through torchvision (idx-ubyte files under ``./data/MNIST/raw``),
resizes every digit to 3000x3000 with PIL on the CPU (0.176 s per image measured) and moves
180 MB of fp32 per batch over PCIe (mnist_onegpu.py:51-66, mnist_distributed.py:69-88).
Here the dataset stays what it is on disk - uint8 28x28 - all the way to the device
(3.9 KB per batch of 5); the resize runs as a CUDA kernel inside the training step
(ops.resize_pil_bilinear, bit-exact with PIL).  There is no network in this environment,
so ``write_synthetic_idx`` produces MNIST-shaped idx files (also usable to run the
unmodified reference offline: torchvision only checks that the files exist).
"""
from __future__ import annotations

import gzip
import os
import struct
from typing import Iterator, Optional, Tuple

import numpy as np
import torch

_FILES = {
    True: ("train-images-idx3-ubyte", "train-labels-idx1-ubyte"),
    False: ("t10k-images-idx3-ubyte", "t10k-labels-idx1-ubyte"),
}


def synthetic_digits(n: int, seed: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """``n`` MNIST-shaped samples: uint8 [n,28,28] and labels uint8 [n].  Each class is a fixed
    smooth template (so a model can actually learn it) modulated per sample."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:28, 0:28].astype(np.float32)
    templates = np.zeros((10, 28, 28), np.float32)
    trng = np.random.default_rng(1234)
    for k in range(10):
        for _ in range(3):
            cy, cx = trng.uniform(6, 22, 2)
            sy, sx = trng.uniform(1.5, 5.0, 2)
            templates[k] += np.exp(-((yy - cy) ** 2 / (2 * sy ** 2) + (xx - cx) ** 2 / (2 * sx ** 2)))
        templates[k] /= templates[k].max()
    templates[:, :4, :] = 0
    templates[:, -4:, :] = 0
    templates[:, :, :4] = 0
    templates[:, :, -4:] = 0           # MNIST's 4-pixel black border
    labels = rng.integers(0, 10, n).astype(np.uint8)
    gain = rng.uniform(0.6, 1.0, (n, 1, 1)).astype(np.float32)
    noise = rng.normal(0, 0.04, (n, 28, 28)).astype(np.float32)
    img = templates[labels] * gain + noise * (templates[labels] > 0.02)
    return np.clip(img * 255.0, 0, 255).astype(np.uint8), labels


def _write_idx(path: str, arr: np.ndarray) -> None:
    with open(path, "wb") as f:
        f.write(struct.pack(">HBB", 0, 8, arr.ndim))
        f.write(struct.pack(">" + "I" * arr.ndim, *arr.shape))
        f.write(arr.astype(np.uint8).tobytes())


def _read_idx(path: str) -> np.ndarray:
    opener = gzip.open if path.endswith(".gz") else open
    with opener(path, "rb") as f:
        zero, dtype, ndim = struct.unpack(">HBB", f.read(4))
        if zero != 0 or dtype != 8:
            raise ValueError(f"{path}: not an unsigned-byte idx file")
        shape = struct.unpack(">" + "I" * ndim, f.read(4 * ndim))
        data = np.frombuffer(f.read(), dtype=np.uint8)
    return data.reshape(shape)


def write_synthetic_idx(root: str = "./data", n_train: int = 60000, n_test: int = 10000, seed: int = 0) -> str:
    """Create ``root/MNIST/raw/*-ubyte`` with synthetic digits.  Returns the raw folder."""
    raw = os.path.join(root, "MNIST", "raw")
    os.makedirs(raw, exist_ok=True)
    for train, n, s in ((True, n_train, seed), (False, n_test, seed + 1)):
        img, lab = synthetic_digits(n, s)
        _write_idx(os.path.join(raw, _FILES[train][0]), img)
        _write_idx(os.path.join(raw, _FILES[train][1]), lab)
    return raw


def raw_files_present(root: str = "./data", train: bool = True) -> bool:
    """True when ``root/MNIST/raw`` holds the idx files (plain or .gz) - the presence-only check torchvision does."""
    raw = os.path.join(root, "MNIST", "raw")
    return all(os.path.exists(os.path.join(raw, f)) or os.path.exists(os.path.join(raw, f) + ".gz")
               for f in _FILES[train])


def open_dataset(root: str = "./data", synthetic: Optional[int] = None, train: bool = True, quiet: bool = False):
    """What the entrypoints train on.  ``synthetic=None`` (the default) reads the real idx files under ``root`` when
    they are there - as the reference does after its download (mnist_onegpu.py:51-54) - and only otherwise falls back
    to 60,000 synthetic digits, saying so; ``synthetic=N`` forces N generated digits."""
    if synthetic is None and raw_files_present(root, train):
        return MNIST(root, train=train)
    n = 60000 if synthetic is None else int(synthetic)
    if synthetic is None and not quiet:
        print(f"note: no MNIST idx files under {os.path.join(root, 'MNIST', 'raw')} (and no network to download "
              f"them): training on {n} synthetic MNIST-shaped digits")
    return MNIST(root, train=train, synthetic=n)


class MNIST:
    """uint8 digits + int64 labels held as CPU tensors.  ``synthetic=N`` skips the disk."""

    def __init__(self, root: str = "./data", train: bool = True, synthetic: Optional[int] = None, seed: int = 0):
        if synthetic is not None:
            img, lab = synthetic_digits(int(synthetic), seed)
        else:
            raw = os.path.join(root, "MNIST", "raw")
            paths = [os.path.join(raw, f) for f in _FILES[train]]
            for i, p in enumerate(paths):
                if not os.path.exists(p) and os.path.exists(p + ".gz"):
                    paths[i] = p + ".gz"
                elif not os.path.exists(p):
                    raise FileNotFoundError(f"{p} not found (no network here: call write_synthetic_idx({root!r}) "
                                            "or pass synthetic=N)")
            img, lab = _read_idx(paths[0]), _read_idx(paths[1])
        self.images = torch.from_numpy(np.ascontiguousarray(img))
        self.labels = torch.from_numpy(lab.astype(np.int64))

    def __len__(self) -> int:
        return self.images.shape[0]

    def __getitem__(self, i):
        return self.images[i], int(self.labels[i])


class DataLoader:
    """Batches of raw digits in pinned memory: ``(uint8 [B,28,28], int64 [B])``.

    Mirrors the reference's loader arguments (batch_size, shuffle, sampler, pin_memory,
    ``drop_last=False``, ``num_workers=0``: mnist_onegpu.py:55-59).  ``shuffle=True`` draws a
    permutation from torch's global RNG like torch's RandomSampler does."""

    def __init__(self, dataset: MNIST, batch_size: int = 5, shuffle: bool = False, sampler=None,
                 pin_memory: bool = True, drop_last: bool = False, num_workers: int = 0):
        if shuffle and sampler is not None:
            raise ValueError("sampler option is mutually exclusive with shuffle")
        self.dataset, self.batch_size, self.shuffle, self.sampler = dataset, int(batch_size), shuffle, sampler
        self.drop_last = drop_last
        self.pin = bool(pin_memory) and torch.cuda.is_available()

    def __len__(self) -> int:
        n = len(self.sampler) if self.sampler is not None else len(self.dataset)
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def _order(self) -> torch.Tensor:
        if self.sampler is not None:
            idx = self.sampler.indices() if hasattr(self.sampler, "indices") else torch.tensor(list(self.sampler))
            return idx
        n = len(self.dataset)
        return torch.randperm(n) if self.shuffle else torch.arange(n)

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        order = self._order()
        B = self.batch_size
        for s in range(0, len(order), B):
            idx = order[s:s + B]
            if len(idx) < B and self.drop_last:
                return
            img = self.dataset.images.index_select(0, idx)
            lab = self.dataset.labels.index_select(0, idx)
            if self.pin:
                img, lab = img.pin_memory(), lab.pin_memory()
            yield img, lab


def get_batch(batch, rank):
    """Move an ``(x, y)`` batch to device ``rank`` when CUDA is present (test_init.py:102-110)."""
    x, y = batch
    if torch.cuda.is_available():
        x, y = x.to(rank, non_blocking=True), y.to(rank, non_blocking=True)
    return x, y
