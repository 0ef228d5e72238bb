"""ctypes loader for the two in-tree native libraries.

* ``libb200dp_host.so``  - C++ runtime: TCP key-value store, shared-memory host
  collectives, DDP bucket planner (replaces torch's C++ TCPStore / ProcessGroupGloo
  / Reducer that the reference reaches through ``dist.init_process_group`` and
  ``DistributedDataParallel``; reference call sites: test_init.py:91,
  mnist_distributed.py:50,67).
* ``libb200dp_cuda.so``  - sm_100a kernels + symmetric-memory runtime.

Both are plain C ABI (see ``include/b200dp.h``), built by ``make`` /
``__graft_entry__.build()`` and kept in-tree so a gpurun snapshot carries them.
A missing CUDA library on a machine that HAS a GPU is a hard error: there is no
silent PyTorch fallback on the hot path.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
_LIBDIR = os.path.join(_HERE, "lib")
_lock = threading.Lock()
_libs: dict[str, ctypes.CDLL] = {}


class NativeError(RuntimeError):
    pass


def build(verbose: bool = False) -> None:
    """Compile both native libraries in-tree (nvcc cross-compiles sm_100a without a GPU)."""
    cmd = ["make", "-C", _ROOT, "-j", str(min(8, os.cpu_count() or 1)), "all"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout[-4000:])
        print(res.stderr[-4000:])
    if res.returncode != 0:
        raise NativeError("native build failed (see output above)")


def _load(name: str) -> ctypes.CDLL:
    with _lock:
        lib = _libs.get(name)
        if lib is not None:
            return lib
        path = os.path.join(_LIBDIR, name)
        if not os.path.exists(path):
            if os.environ.get("B200DP_NO_AUTOBUILD") == "1":
                raise NativeError(f"{path} is missing; run `make` at the repo root")
            build()
        try:
            lib = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
        except OSError as e:  # pragma: no cover - depends on the box
            raise NativeError(f"cannot load {path}: {e}") from e
        _libs[name] = lib
        return lib


def host() -> ctypes.CDLL:
    return _load("libb200dp_host.so")


def cuda() -> ctypes.CDLL:
    lib = _load("libb200dp_cuda.so")
    if not getattr(lib, "_b200_bound", False):
        lib.b200_last_error.restype = ctypes.c_char_p
        lib._b200_bound = True
    return lib


def check(rc: int, what: str) -> None:
    """Raise with the native error message if a C-ABI call failed."""
    if rc != 0:
        msg = cuda().b200_last_error()
        raise NativeError(f"{what} failed (rc={rc}): {msg.decode() if msg else '?'}")


def host_check(rc: int, what: str) -> None:
    if rc != 0:
        lib = host()
        lib.b200h_last_error.restype = ctypes.c_char_p
        msg = lib.b200h_last_error()
        raise NativeError(f"{what} failed (rc={rc}): {msg.decode() if msg else '?'}")
