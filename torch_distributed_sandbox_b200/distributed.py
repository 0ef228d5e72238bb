"""Process groups and collectives: the ``torch.distributed`` surface the reference uses,
re-implemented on the b200dp runtime (L2/L3, D2-D6 in SURVEY.md).

Reference call sites covered:
  * ``dist.init_process_group(backend, rank=, world_size=)`` with the implicit ``env://``
    method                                  test_init.py:91, allreduce_toy.py:44, mnist_distributed.py:50
  * ``dist.destroy_process_group()``        test_init.py:96-100
  * ``dist.new_group(ranks=...)`` per step  allreduce_toy.py:27, mnist_distributed.py:100
  * ``dist.all_reduce(t, op=SUM, group=)``  allreduce_toy.py:31 (and AVG, commented at mnist_distributed.py:102)
  * ``dist.barrier()``                      allreduce_toy.py:33
  * broadcast / all_gather (used implicitly by DDP's constructor and buffer sync)

Backends
  ``"nccl"`` / ``"native"``  GPU tensors; collectives are our own sm_100a kernels over
                             symmetric (peer-mapped / multicast) memory - no NCCL call.
                             The name ``nccl`` is accepted because the reference hard-codes it.
  ``"gloo"`` / ``"host"``    CPU tensors; shared-memory collectives (csrc/hostcoll.cpp).
``backend=None`` picks native when CUDA is available, else host (test_init.py:84-88).

Unlike torch, ``new_group`` caches by rank tuple: the reference builds a fresh group every
training step and leaks a communicator each time; here that is a dictionary lookup (C7).
"""
from __future__ import annotations

import atexit
import ctypes
import enum
import os
import secrets
import time
from ctypes import c_char_p, c_int, c_size_t, c_void_p
from typing import Dict, List, Optional, Sequence, Tuple
from urllib.parse import parse_qs, urlparse

import torch

from . import _native
from .store import StoreError, StoreTimeout, TCPStore

DEFAULT_TIMEOUT_S = 600.0


class ReduceOp(enum.IntEnum):
    SUM = 0
    AVG = 1
    MAX = 2
    MIN = 3
    PRODUCT = 4


class Work:
    """Handle returned by ``async_op=True`` collectives."""

    def __init__(self, event: Optional["torch.cuda.Event"] = None, stream=None):
        self._event = event
        self._stream = stream

    def wait(self) -> bool:
        if self._event is not None:
            torch.cuda.current_stream().wait_event(self._event)
        return True

    def is_completed(self) -> bool:
        return True if self._event is None else self._event.query()


# --------------------------------------------------------------------- backends --
_DT = {torch.float32: 0, torch.float64: 1, torch.int32: 2, torch.int64: 3, torch.uint8: 4, torch.bfloat16: 5}


def _hlib():
    lib = _native.host()
    if not getattr(lib, "_b200_coll_bound", False):
        lib.b200h_coll_last_error.restype = c_char_p
        lib.b200h_coll_open.restype = c_void_p
        lib.b200h_coll_open.argtypes = [c_char_p, c_int, c_int, c_int, c_size_t, c_int]
        lib.b200h_coll_close.argtypes = [c_void_p, c_int]
        lib.b200h_coll_barrier.argtypes = [c_void_p]
        lib.b200h_coll_allreduce.argtypes = [c_void_p, c_void_p, c_size_t, c_int, c_int]
        lib.b200h_coll_broadcast.argtypes = [c_void_p, c_void_p, c_size_t, c_int]
        lib.b200h_coll_allgather.argtypes = [c_void_p, c_void_p, c_void_p, c_size_t]
        lib._b200_coll_bound = True
    return lib


class DistError(RuntimeError):
    pass


class HostBackend:
    """Shared-memory collectives for CPU tensors (single node)."""

    name = "host"
    SLOT_BYTES = 4 << 20

    def __init__(self, group_rank: int, size: int, seg_name: str, timeout_s: float):
        self.rank, self.size = group_rank, size
        self._name = seg_name
        self._h = _hlib().b200h_coll_open(seg_name.encode(), group_rank, size, 1 if group_rank == 0 else 0,
                                          self.SLOT_BYTES, int(timeout_s * 1000))
        if not self._h:
            raise DistError(_hlib().b200h_coll_last_error().decode())

    def _chk(self, rc):
        if rc != 0:
            raise DistError(_hlib().b200h_coll_last_error().decode())

    @staticmethod
    def _stage(t: torch.Tensor) -> Tuple[torch.Tensor, bool]:
        if t.device.type != "cpu" or not t.is_contiguous():
            return t.detach().to("cpu").contiguous(), True
        return t, False

    def all_reduce(self, t: torch.Tensor, op: ReduceOp):
        if t.dtype not in _DT:
            raise DistError(f"all_reduce: dtype {t.dtype} is not supported by the host backend")
        buf, staged = self._stage(t)
        self._chk(_hlib().b200h_coll_allreduce(self._h, buf.data_ptr(), buf.numel(), _DT[t.dtype], int(op)))
        if staged:
            t.copy_(buf)

    def broadcast(self, t: torch.Tensor, src: int):
        buf, staged = self._stage(t)
        self._chk(_hlib().b200h_coll_broadcast(self._h, buf.data_ptr(), buf.numel() * buf.element_size(), src))
        if staged:
            t.copy_(buf)

    def all_gather(self, out: torch.Tensor, t: torch.Tensor):
        src, _ = self._stage(t)
        dst = out if (out.device.type == "cpu" and out.is_contiguous()) else torch.empty(out.shape, dtype=out.dtype)
        self._chk(_hlib().b200h_coll_allgather(self._h, src.data_ptr(), dst.data_ptr(),
                                               src.numel() * src.element_size()))
        if dst is not out:
            out.copy_(dst)

    def barrier(self):
        self._chk(_hlib().b200h_coll_barrier(self._h))

    def close(self):
        if self._h:
            _hlib().b200h_coll_close(self._h, 1 if self.rank == 0 else 0)
            self._h = None


class ProcessGroup:
    def __init__(self, gid: int, ranks: Sequence[int], my_global_rank: int, backend):
        self.id = gid
        self.ranks = tuple(ranks)
        self._rank = self.ranks.index(my_global_rank)
        self.backend = backend

    def rank(self) -> int:
        return self._rank

    def size(self) -> int:
        return len(self.ranks)

    def __repr__(self):
        return f"ProcessGroup(id={self.id}, ranks={self.ranks}, backend={self.backend.name})"


class _NonMember:
    def __repr__(self):
        return "GroupMember.NON_MEMBER"


class GroupMember:
    NON_MEMBER = _NonMember()


class _World:
    def __init__(self):
        self.reset()

    def reset(self):
        self.store: Optional[TCPStore] = None
        self.rank = -1
        self.world_size = -1
        self.backend_name = ""
        self.default: Optional[ProcessGroup] = None
        self.groups: Dict[Tuple[int, ...], object] = {}
        self.group_count = 0
        self.job = ""
        self.timeout = DEFAULT_TIMEOUT_S
        self.device: Optional[torch.device] = None
        self.init_seconds = 0.0


_W = _World()


# ------------------------------------------------------------------- rendezvous --
def _parse_rendezvous(init_method: Optional[str], rank: int, world_size: int):
    """Resolve (master_addr, master_port, rank, world_size) from ``env://`` (default, as the
    reference relies on) or ``tcp://host:port[?rank=..&world_size=..]`` with the sanity checks
    torch's env handler performs: every missing piece is named in the error."""
    init_method = init_method or "env://"
    if init_method.startswith("env://"):
        missing = [k for k in ("MASTER_ADDR", "MASTER_PORT") if not os.environ.get(k)]
        if rank < 0:
            if os.environ.get("RANK") is None:
                missing.append("RANK")
            else:
                rank = int(os.environ["RANK"])
        if world_size < 0:
            if os.environ.get("WORLD_SIZE") is None:
                missing.append("WORLD_SIZE")
            else:
                world_size = int(os.environ["WORLD_SIZE"])
        if missing:
            raise ValueError("env:// rendezvous: environment variable(s) " + ", ".join(missing) +
                             " expected but not set (pass rank=/world_size= or export them)")
        addr, port = os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"]
    elif init_method.startswith("tcp://"):
        u = urlparse(init_method)
        if not u.hostname or not u.port:
            raise ValueError(f"tcp:// rendezvous needs host and port, got {init_method!r}")
        q = parse_qs(u.query)
        if rank < 0 and "rank" in q:
            rank = int(q["rank"][0])
        if world_size < 0 and "world_size" in q:
            world_size = int(q["world_size"][0])
        if rank < 0 or world_size < 0:
            raise ValueError("tcp:// rendezvous: rank and world_size must be given (arguments or query string)")
        addr, port = u.hostname, str(u.port)
    else:
        raise ValueError(f"unsupported init_method {init_method!r} (use env:// or tcp://host:port)")
    try:
        port_i = int(port)
    except ValueError:
        raise ValueError(f"MASTER_PORT must be an integer, got {port!r}") from None
    if not (0 < port_i < 65536):
        raise ValueError(f"MASTER_PORT {port_i} out of range")
    if world_size < 1:
        raise ValueError(f"world_size must be >= 1, got {world_size}")
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} is outside [0, world_size={world_size})")
    if addr == "localhost":
        addr = "127.0.0.1"
    return addr, port_i, rank, world_size


def _bootstrap_store(addr: str, port: int, rank: int, world_size: int, timeout: float) -> TCPStore:
    """Rank 0 hosts our store.  Under ``torchrun`` the agent already owns MASTER_PORT with a
    torch TCPStore (TORCHELASTIC_USE_AGENT_STORE=True): then rank 0 binds a free port and
    publishes it through the agent's store, which is used for nothing else."""
    if os.environ.get("TORCHELASTIC_USE_AGENT_STORE", "").lower() == "true":
        from datetime import timedelta
        from torch.distributed import TCPStore as _AgentStore
        agent = _AgentStore(addr, port, world_size, False, timedelta(seconds=timeout))
        key = f"b200dp/store_port/{os.environ.get('TORCHELASTIC_RUN_ID', '0')}/{os.environ.get('TORCHELASTIC_RESTART_COUNT', '0')}"
        if rank == 0:
            st = TCPStore(addr, 0, True, timeout)
            agent.set(key, str(st.port))
            return st
        real_port = int(agent.get(key).decode())
        return TCPStore(addr, real_port, False, timeout)
    return TCPStore(addr, port, rank == 0, timeout)


def _make_backend(kind: str, gid: int, ranks: Sequence[int]):
    grank = list(ranks).index(_W.rank)
    if kind == "host":
        return HostBackend(grank, len(ranks), f"/b200dp_{_W.job}_{gid}", _W.timeout)
    from .parallel.symm_backend import SymmBackend
    return SymmBackend(_W.store, _W.job, gid, tuple(ranks), _W.rank, _W.device, _W.timeout)


def _canon_backend(backend: Optional[str]) -> str:
    if backend is None:
        return "native" if torch.cuda.is_available() else "host"
    b = str(backend).lower()
    if b in ("nccl", "native", "b200", "symm"):
        if not torch.cuda.is_available():
            raise RuntimeError(f"backend {backend!r} needs CUDA devices; use 'gloo' on a CPU-only machine")
        return "native"
    if b in ("gloo", "host", "shm", "cpu"):
        return "host"
    raise ValueError(f"unknown backend {backend!r} (expected nccl/native or gloo/host)")


def init_process_group(backend: Optional[str] = None, init_method: Optional[str] = None,
                       timeout: Optional[float] = None, world_size: int = -1, rank: int = -1,
                       device_id=None) -> None:
    """Bring up the default process group (test_init.py:91 / mnist_distributed.py:50 semantics)."""
    if _W.default is not None:
        raise RuntimeError("trying to initialize the default process group twice")
    t0 = time.time()
    if hasattr(timeout, "total_seconds"):
        timeout = timeout.total_seconds()
    timeout = float(timeout) if timeout is not None else DEFAULT_TIMEOUT_S
    kind = _canon_backend(backend)
    addr, port, rank, world_size = _parse_rendezvous(init_method, rank, world_size)
    if device_id is not None:
        dev = torch.device(device_id) if not isinstance(device_id, int) else torch.device("cuda", device_id)
        if dev.type != "cuda" or dev.index is None:
            raise ValueError(f"device_id must be a CUDA device with an explicit index, got {device_id!r}")
        if dev.index >= torch.cuda.device_count():
            raise ValueError(f"device_id {dev} is out of range ({torch.cuda.device_count()} devices visible)")
        _W.device = dev
    store = _bootstrap_store(addr, port, rank, world_size, timeout)
    try:
        if store.add(f"rank_claim/{rank}", 1) != 1:
            raise RuntimeError(f"rank {rank} was claimed by more than one process (duplicate rank)")
        if rank == 0:
            store.set("job/token", f"{port}_{secrets.token_hex(4)}")
            store.set("job/world_size", str(world_size))
        try:
            ws_master = int(store.get("job/world_size", timeout).decode())
        except StoreTimeout as e:
            raise RuntimeError(f"rendezvous timed out: rank 0 never published the job ({e})") from None
        if ws_master != world_size:
            raise RuntimeError(f"world_size mismatch: this rank was started with {world_size}, rank 0 with {ws_master}")
        job = store.get("job/token", timeout).decode()
        n = store.add("job/arrived", 1)
        if n == world_size:
            store.set("job/all_arrived", b"1")
        try:
            store.get("job/all_arrived", timeout)
        except StoreTimeout:
            have = store.add("job/arrived", 0)
            raise RuntimeError(f"rendezvous timed out after {timeout:.0f}s: only {have} of {world_size} ranks "
                               f"reached {addr}:{port}") from None
    except BaseException:
        store.close()
        raise
    _W.store, _W.rank, _W.world_size, _W.backend_name, _W.job, _W.timeout = store, rank, world_size, kind, job, timeout
    ranks = tuple(range(world_size))
    _W.default = ProcessGroup(0, ranks, rank, _make_backend(kind, 0, ranks))
    _W.groups[ranks] = _W.default
    _W.group_count = 1
    _W.init_seconds = time.time() - t0


def is_initialized() -> bool:
    return _W.default is not None


def is_available() -> bool:
    return True


def _default() -> ProcessGroup:
    if _W.default is None:
        raise RuntimeError("default process group has not been initialized; call init_process_group first")
    return _W.default


def get_rank(group=None) -> int:
    if isinstance(group, _NonMember):
        return -1
    return (group or _default()).rank()


def get_world_size(group=None) -> int:
    if isinstance(group, _NonMember):
        return -1
    return (group or _default()).size()


def get_backend(group=None) -> str:
    return (group or _default()).backend.name


def get_store() -> TCPStore:
    _default()
    return _W.store


def new_group(ranks: Optional[Sequence[int]] = None, backend: Optional[str] = None):
    """Collective over ALL ranks of the job (like torch): returns the group for members and
    ``GroupMember.NON_MEMBER`` otherwise.  Same rank tuple -> same cached group, no new
    handshake (the reference calls this every step: allreduce_toy.py:26-27)."""
    _default()
    ranks = tuple(sorted(set(range(_W.world_size) if ranks is None else (int(r) for r in ranks))))
    if not ranks or ranks[0] < 0 or ranks[-1] >= _W.world_size:
        raise ValueError(f"new_group ranks {ranks} outside the world of {_W.world_size}")
    if ranks in _W.groups:
        return _W.groups[ranks]
    gid = _W.group_count
    _W.group_count += 1
    if _W.rank not in ranks:
        _W.groups[ranks] = GroupMember.NON_MEMBER
        return GroupMember.NON_MEMBER
    kind = _canon_backend(backend) if backend is not None else _W.backend_name
    g = ProcessGroup(gid, ranks, _W.rank, _make_backend(kind, gid, ranks))
    _W.groups[ranks] = g
    return g


def _resolve(group) -> Optional[ProcessGroup]:
    if isinstance(group, _NonMember):
        return None
    return group or _default()


def all_reduce(tensor: torch.Tensor, op: ReduceOp = ReduceOp.SUM, group=None, async_op: bool = False):
    """In-place all-reduce (allreduce_toy.py:31).  0-dim and integer tensors are fine."""
    g = _resolve(group)
    if g is None:
        return None
    w = g.backend.all_reduce(tensor, ReduceOp(int(op)))
    return (w or Work()) if async_op else None


def broadcast(tensor: torch.Tensor, src: int = 0, group=None, async_op: bool = False):
    """``src`` is a GLOBAL rank, as in torch."""
    g = _resolve(group)
    if g is None:
        return None
    if src not in g.ranks:
        raise ValueError(f"broadcast source rank {src} is not in {g}")
    w = g.backend.broadcast(tensor, g.ranks.index(src))
    return (w or Work()) if async_op else None


def all_gather_into_tensor(output: torch.Tensor, tensor: torch.Tensor, group=None):
    g = _resolve(group)
    if g is None:
        return None
    if output.numel() != tensor.numel() * g.size():
        raise ValueError("all_gather_into_tensor: output must hold world_size * input elements")
    g.backend.all_gather(output, tensor)


def all_gather(tensor_list: List[torch.Tensor], tensor: torch.Tensor, group=None):
    g = _resolve(group)
    if g is None:
        return None
    out = torch.empty((g.size(),) + tuple(tensor.shape), dtype=tensor.dtype, device=tensor.device)
    g.backend.all_gather(out, tensor)
    for i, t in enumerate(tensor_list):
        t.copy_(out[i])


def barrier(group=None, async_op: bool = False):
    """allreduce_toy.py:33.  On the native backend this is a flag barrier in symmetric memory
    plus a stream sync; on the host backend a sense-reversing shared-memory barrier."""
    g = _resolve(group)
    if g is None:
        return None
    g.backend.barrier()
    return Work() if async_op else None


def destroy_process_group(group=None) -> None:
    """test_init.py:96-100.  Destroying the default group tears everything down; rank 0 keeps
    the store alive until every rank has checked out."""
    if _W.default is None:
        return
    if group is not None and group is not _W.default and not isinstance(group, _NonMember):
        group.backend.close()
        _W.groups = {k: v for k, v in _W.groups.items() if v is not group}
        return
    store, rank, world = _W.store, _W.rank, _W.world_size
    for g in list(_W.groups.values()):
        if isinstance(g, ProcessGroup):
            try:
                g.backend.close()
            except Exception:
                pass
    try:
        store.add("job/departed", 1)
        if rank == 0:
            deadline = time.time() + min(_W.timeout, 30.0)
            while store.add("job/departed", 0) < world and time.time() < deadline:
                time.sleep(0.005)
    except StoreError:
        pass
    store.close()
    _W.reset()


@atexit.register
def _atexit_cleanup():  # the reference never calls destroy (test_init.py:96-100 is dead code)
    if _W.default is not None:
        try:
            destroy_process_group()
        except Exception:
            pass
