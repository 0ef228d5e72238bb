"""Native GPU backend: collectives as our own sm_100a kernels over symmetric memory.

This is what stands where ``ProcessGroupNCCL`` stands under the reference
(``init_process_group('nccl')``, mnist_distributed.py:50): every rank maps every peer's
buffers (CUDA VMM + fd passing, csrc/symm.cu), binds them to an NVLS multicast object when
the platform allows, and the collectives are kernels that read/write peer memory over
NVLink (csrc/allreduce.cu).  No NCCL call is made on any path.

Buffers
  * ``alloc_symmetric(numel, dtype)`` - collective; returns a torch tensor living in
    symmetric memory.  ``all_reduce`` on (a slice of) such a tensor runs IN PLACE with the
    two-shot NVLS / P2P kernel: this is how DDP gradient buckets avoid any staging copy.
  * every other tensor goes through a 32 MiB symmetric staging buffer (one-shot kernel for
    small messages, chunked two-shot for large ones).
Flag channels: 0 = collectives issued on the caller's stream, 1 = the DDP side stream, so the
overlapped bucket all-reduce never shares flag words with a concurrent small collective.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_size_t, c_uint32, c_void_p
from typing import Dict, List, Optional, Tuple

import torch

from .. import _native

ONESHOT_MAX_BYTES = 256 << 10
PUSH_SLOT_BYTES = 512 << 10           # per-rank slot of the push ring (small-message all-reduce)
PUSH_MAX_RANKS = 16
STAGING_BYTES = 32 << 20
_DT = {torch.float32: 0, torch.float64: 1, torch.int32: 2, torch.int64: 3, torch.bfloat16: 5}


def _lib():
    lib = _native.cuda()
    if not getattr(lib, "_b200_symm_bound", False):
        lib.b200_symm_comm_create.restype = c_void_p
        lib.b200_symm_comm_create.argtypes = [c_char_p, c_int, c_int, c_int, c_int]
        lib.b200_symm_comm_destroy.argtypes = [c_void_p]
        lib.b200_symm_alloc.restype = c_void_p
        lib.b200_symm_alloc.argtypes = [c_void_p, c_size_t, c_int, ctypes.POINTER(c_void_p), ctypes.POINTER(c_void_p),
                                        ctypes.POINTER(c_size_t), ctypes.POINTER(c_void_p), c_int]
        lib.b200_symm_free.argtypes = [c_void_p, c_void_p]
        lib.b200_symm_topology.argtypes = [c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_int), c_int]
        lib.b200_ar_flag_bytes.restype = c_size_t
        PP = ctypes.POINTER(c_void_p)
        lib.b200_ar_barrier.argtypes = [PP, c_int, c_int, c_uint32, c_void_p]
        lib.b200_ar_oneshot.argtypes = [PP, PP, c_int, c_int, c_void_p, c_size_t, c_int, c_int, c_float, c_int,
                                        c_uint32, c_void_p]
        lib.b200_ar_oneshot_f32x4.argtypes = [PP, PP, c_int, c_int, c_void_p, c_void_p, c_size_t, c_float, c_float, c_int,
                                              c_uint32, c_void_p]
        lib.b200_ar_large_f32.argtypes = [PP, c_void_p, PP, PP, c_void_p, c_int, c_int, c_size_t, c_float, c_float,
                                          c_int, c_int, c_uint32, c_void_p]
        lib.b200_ar_nvls_bf16.argtypes = [c_void_p, PP, c_int, c_int, c_size_t, c_float, c_int, c_uint32, c_void_p]
        lib.b200_ar_nvls_bf16_sgd.argtypes = [c_void_p, PP, PP, c_void_p, c_int, c_int, c_size_t, c_float, c_float,
                                              c_int, c_uint32, c_void_p]
        lib.b200_ar_broadcast.argtypes = [PP, PP, c_int, c_int, c_void_p, c_size_t, c_int, c_int, c_uint32, c_void_p]
        lib.b200_ar_allgather.argtypes = [PP, PP, c_int, c_int, c_void_p, c_size_t, c_int, c_uint32, c_void_p]
        lib._b200_symm_bound = True
    return lib


def _on_device(fn):
    """Run a collective with the group's device current (so `current_stream()` is that device's stream) -
    a no-op in the usual one-GPU-per-process set-up where it already is."""
    import functools

    @functools.wraps(fn)
    def wrapper(self, *a, **k):
        if self.size > 1 and (self._comm is None or self.device is None):
            self._ensure(next((x for x in a if isinstance(x, torch.Tensor) and x.is_cuda), None))
        dev = self.device
        if dev is None or self.size == 1 or torch.cuda.current_device() == dev.index:
            return fn(self, *a, **k)
        with torch.cuda.device(dev):
            return fn(self, *a, **k)
    return wrapper


class _CudaView:
    """Minimal ``__cuda_array_interface__`` carrier so torch can wrap raw symmetric memory."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}
        self._owner = owner


class SymmBuffer:
    def __init__(self, backend: "SymmBackend", nbytes: int, multicast: bool = True):
        lib = _lib()
        world = backend.size
        ptrs = (c_void_p * 16)()
        mc, padded, table = c_void_p(0), c_size_t(0), c_void_p(0)
        self._backend = backend
        self._h = lib.b200_symm_alloc(backend._comm, nbytes, 1 if multicast else 0, ptrs, ctypes.byref(mc),
                                      ctypes.byref(padded), ctypes.byref(table), int(backend.timeout * 1000))
        if not self._h:
            raise RuntimeError("symmetric allocation failed: " + _native.cuda().b200_last_error().decode())
        self.ptrs: List[int] = [int(ptrs[i] or 0) for i in range(world)]
        self.mc: int = int(mc.value or 0)
        self.nbytes = int(padded.value)
        self.local = self.ptrs[backend.rank]
        with torch.cuda.device(backend.device):
            self.bytes_view = torch.as_tensor(_CudaView(self.local, self.nbytes, self), device=backend.device)

    def tensor(self, numel: int, dtype: torch.dtype, byte_offset: int = 0) -> torch.Tensor:
        es = torch.empty((), dtype=dtype).element_size()
        return self.bytes_view[byte_offset: byte_offset + numel * es].view(dtype)

    def ptr_array(self, byte_offset: int = 0):
        arr = (c_void_p * 16)()
        for i, p in enumerate(self.ptrs):
            arr[i] = p + byte_offset
        return arr

    def free(self):
        if self._h:
            _lib().b200_symm_free(self._backend._comm, self._h)
            self._h = None


class SymmBackend:
    name = "native"

    def __init__(self, store, job: str, gid: int, ranks: Tuple[int, ...], my_global_rank: int,
                 device: Optional[torch.device], timeout: float):
        self.ranks, self.size = tuple(ranks), len(ranks)
        self.rank = self.ranks.index(my_global_rank)
        self.timeout = timeout
        self._buffers: List[SymmBuffer] = []
        self._comm = None
        self.epoch = [4, 4]
        self.multicast = False
        self._tag = f"b200dp_{job}_{gid}".encode()
        # Like ProcessGroupNCCL, the communicator is bound to a device LAZILY: the reference calls
        # init_process_group before torch.cuda.set_device (mnist_distributed.py:50,57) and allreduce_toy.py never
        # calls set_device at all, so the device is only known at the first collective (from its tensor, or the
        # then-current device).  With an explicit device_id the bring-up happens right away.
        self.device: Optional[torch.device] = torch.device(device) if device is not None else None
        if self.size > 1 and self.device is not None:
            self._bring_up()

    def _ensure(self, t: Optional[torch.Tensor] = None) -> None:
        """COLLECTIVE on first use: bind to the tensor's device (else the current device) and map the peers."""
        if self._comm is not None or self.size == 1:
            if self.device is None:
                self.device = t.device if (t is not None and t.is_cuda) else torch.device("cuda", torch.cuda.current_device())
            if t is not None and t.is_cuda and t.device != self.device:
                raise RuntimeError(f"collective on a tensor of {t.device}, but this process group is bound to "
                                   f"{self.device} (one process drives one GPU; pass device_id= or call "
                                   f"torch.cuda.set_device before the first collective)")
            return
        if self.device is None:
            self.device = t.device if (t is not None and t.is_cuda) else torch.device("cuda", torch.cuda.current_device())
        self._bring_up()

    def _bring_up(self) -> None:
        lib = _lib()
        timeout = self.timeout
        with torch.cuda.device(self.device):
            torch.zeros(1, device=self.device)          # make sure the primary context exists
        prev = torch.cuda.current_device()
        torch.cuda.set_device(self.device)
        try:
            self._bring_up_locked(lib, timeout)
        finally:
            torch.cuda.set_device(prev)

    def _bring_up_locked(self, lib, timeout) -> None:
        tag = self._tag
        self._comm = lib.b200_symm_comm_create(tag, self.rank, self.size, self.device.index, int(timeout * 1000))
        if not self._comm:
            raise RuntimeError("symmetric-memory bring-up failed: " + _native.cuda().b200_last_error().decode())
        lib.b200_ar_set_timeout_ms.argtypes = [ctypes.c_longlong]
        _native.check(lib.b200_ar_set_timeout_ms(max(1, int(timeout * 1000))), "set collective timeout")
        fb = int(lib.b200_ar_flag_bytes())
        self._flags = SymmBuffer(self, 2 * fb, multicast=False)
        self._flag_arr = [self._flags.ptr_array(0), self._flags.ptr_array(fb)]
        self._staging = SymmBuffer(self, STAGING_BYTES, multicast=True)
        # push ring for small messages: [2 halves][16 slots][512 KiB], one copy on every rank
        self._push = SymmBuffer(self, 2 * PUSH_MAX_RANKS * PUSH_SLOT_BYTES, multicast=False)
        self._push_calls = 0
        lib.b200_ar_push.argtypes = [ctypes.POINTER(c_void_p), ctypes.POINTER(c_void_p), c_int, c_int, c_void_p, c_void_p,
                                     c_size_t, c_int, c_int, c_float, c_size_t, c_int, c_int, c_uint32, c_void_p]
        self._push_arr = self._push.ptr_array(0)
        self._buffers += [self._flags, self._staging, self._push]
        self.multicast = self._staging.mc != 0
        self.comm_stream = torch.cuda.Stream(device=self.device, priority=-1)
        self.oneshot_max_bytes = self.algo_policy()["oneshot_max_bytes"]
        if self.rank == 0 and os.environ.get("B200DP_QUIET", "0") != "1":
            t = self.topology()
            print(f"[b200dp] process group up: world {self.size}, {self.device}: NVLink {t['nvlink_links_active']} links "
                  f"active ({t['nvlink_links_to_switch']} to NVSwitch, {t['nvlink_link_gbs']} GB/s each), "
                  f"multicast {'bound' if self.multicast else 'not bound'}; "
                  f"all-reduce: one-shot <= {self.oneshot_max_bytes >> 10} KiB, then {self.algo_policy()['large']}",
                  flush=True)

    # ------------------------------------------------------------- helpers --
    def describe(self) -> Dict:
        return {"backend": "native-symmetric", "world": self.size, "nvls_multicast": bool(self.multicast)}

    def topology(self) -> Dict:
        """Fabric report for this rank's GPU: CUDA peer access plus the NVML view of its NVLink links (how many are
        up, how many end in an NVSwitch, direct links per peer, per-link speed) and NVLS multicast support."""
        n = torch.cuda.device_count()
        p2p, at = (c_int * 64)(), (c_int * 64)()
        lib = _lib()
        dev = self.device.index if self.device is not None else torch.cuda.current_device()
        cnt = lib.b200_symm_topology(dev, p2p, at, 64)
        out = {"devices": cnt, "p2p": [int(p2p[i]) for i in range(min(cnt, n))],
               "native_atomics": [int(at[i]) for i in range(min(cnt, n))], "multicast": bool(self.multicast)}
        la, ls, gbs, mc = c_int(0), c_int(0), c_int(0), c_int(0)
        peer = (c_int * 64)()
        IP = ctypes.POINTER(c_int)
        lib.b200_nvlink_report.argtypes = [c_int, IP, IP, IP, c_int, IP, IP]
        rc = lib.b200_nvlink_report(dev, ctypes.byref(la), ctypes.byref(ls), peer, 64, ctypes.byref(gbs), ctypes.byref(mc))
        out.update({"nvml": rc == 0, "nvlink_links_active": la.value, "nvlink_links_to_switch": ls.value,
                    "nvlink_direct_links": [int(peer[i]) for i in range(min(cnt, n))], "nvlink_link_gbs": gbs.value,
                    "nvswitch": ls.value > 0, "multicast_supported": bool(mc.value)})
        return out

    def algo_policy(self) -> Dict:
        """Collective algorithm thresholds chosen from the fabric and the world size (measured on 2/4/8 x B200,
        profiles/allreduce_sweep_*): below `oneshot_max_bytes` every rank reads every peer (latency-bound, (n-1) x S
        of inbound traffic is still cheap); above it the bucket is sharded - through the switch's reduction
        (multimem) when an NVSwitch with multicast is there, with peer loads otherwise."""
        t = self.topology()
        nvls = bool(self.multicast and t["multicast_supported"])
        n = self.size
        oneshot = (512 << 10) if n <= 2 else (256 << 10) if n <= 4 else (128 << 10)
        return {"large": "nvls" if nvls else "p2p", "oneshot_max_bytes": oneshot,
                "why": ("NVSwitch + multicast: in-switch reduction" if nvls else
                        "no multicast binding: two-shot over peer loads") + f"; world {n}"}

    def _next_epoch(self, channel: int) -> int:
        e = self.epoch[channel]
        self.epoch[channel] = (e + 4) & 0xFFFFFFFF
        return e

    @staticmethod
    def _stream() -> c_void_p:
        return c_void_p(torch.cuda.current_stream().cuda_stream)

    def alloc_symmetric(self, numel: int, dtype: torch.dtype = torch.float32, multicast: bool = True) -> torch.Tensor:
        """COLLECTIVE.  A zero-initialised tensor in symmetric (peer-mapped, multicast-bound) memory."""
        es = torch.empty((), dtype=dtype).element_size()
        self._ensure()
        if self.size == 1:
            return torch.zeros(numel, dtype=dtype, device=self.device)
        with torch.cuda.device(self.device):
            buf = SymmBuffer(self, max(numel * es, 16), multicast=multicast)
        self._buffers.append(buf)
        return buf.tensor(numel, dtype)

    def free_symmetric(self, t: torch.Tensor) -> bool:
        """COLLECTIVE.  Release the symmetric buffer `t` was allocated from (peer mappings, multicast binding and the
        physical memory); every rank must call it for the same buffer.  The tensor must not be used afterwards."""
        hit = self.lookup(t) if (self.size > 1 and t.is_cuda) else None
        if hit is None or hit[0] in (getattr(self, "_flags", None), getattr(self, "_staging", None),
                                     getattr(self, "_push", None)):
            return False
        torch.cuda.synchronize(self.device)
        self.barrier()                       # no peer kernel may still be reading the buffer
        self._buffers.remove(hit[0])
        hit[0].free()
        return True

    def lookup(self, t: torch.Tensor) -> Optional[Tuple[SymmBuffer, int]]:
        p = t.data_ptr()
        for b in self._buffers:
            if b.local <= p < b.local + b.nbytes:
                return b, p - b.local
        return None

    def _blocks(self, nbytes: int) -> int:
        return max(1, min(64, (nbytes + (1 << 16) - 1) >> 16))

    # --------------------------------------------------------- collectives --
    @_on_device
    def all_reduce_bf16_sgd(self, grad_bf16: torch.Tensor, weights: torch.Tensor, lr: float, scale: float,
                            channel: int = 1, blocks: int = 8):
        """DDP bucket exchanged in bf16: ``weights -= lr * scale * sum_ranks(grad_bf16)`` on every rank in one pass
        (switch-side fp32 accumulation of the bf16 values, SGD on the fp32 master weights, new weights multicast).
        Both tensors must be symmetric with a multicast mapping (NVLS)."""
        if self.size == 1:
            weights.add_(grad_bf16.view_as(weights).float(), alpha=-float(lr) * scale)
            return
        self._ensure(grad_bf16)
        g, w = self.lookup(grad_bf16.view(-1)), self.lookup(weights.view(-1))
        if g is None or w is None or not g[0].mc or not w[0].mc:
            raise RuntimeError("all_reduce_bf16_sgd needs both buckets in symmetric memory with NVLS multicast")
        assert grad_bf16.dtype == torch.bfloat16 and weights.dtype == torch.float32
        assert grad_bf16.numel() == weights.numel() and grad_bf16.numel() % 8 == 0 and g[1] % 16 == 0 and w[1] % 16 == 0
        rc = _lib().b200_ar_nvls_bf16_sgd(g[0].mc + g[1], self._flag_arr[channel], w[0].ptr_array(w[1]), w[0].mc + w[1],
                                          self.rank, self.size, grad_bf16.numel(), float(scale), float(lr), int(blocks),
                                          self._next_epoch(channel), self._stream())
        _native.check(rc, "all_reduce_bf16_sgd")

    @_on_device
    def all_reduce(self, t: torch.Tensor, op, channel: int = 0, sgd: Optional[Tuple[torch.Tensor, float]] = None,
                   scale: Optional[float] = None, algo: Optional[str] = None, blocks: Optional[int] = None):
        """In-place all-reduce.  ``sgd=(weights, lr)`` fuses ``weights -= lr * result`` into the
        reduction epilogue (both tensors must be symmetric, fp32, same shape)."""
        op = int(op)
        if self.size == 1:
            if sgd is not None:
                sgd[0].add_(t.view_as(sgd[0]), alpha=-float(sgd[1]) * (scale or 1.0))
            return None
        if t.dtype not in _DT:
            raise RuntimeError(f"all_reduce: dtype {t.dtype} not supported by the native backend")
        if not t.is_cuda:
            raise RuntimeError("all_reduce on the native backend needs a CUDA tensor (use backend='gloo' for CPU)")
        self._ensure(t)
        if scale is None:
            scale = 1.0 / self.size if op == 1 else 1.0
        if op == 1:
            if not t.dtype.is_floating_point:
                raise RuntimeError("ReduceOp.AVG needs a floating-point tensor")
            op = 0
        lib, es = _lib(), t.element_size()
        nbytes = t.numel() * es
        flat = t.contiguous().view(-1) if not t.is_contiguous() else t.view(-1)
        sym = self.lookup(flat)
        large_ok = t.dtype == torch.float32 and op == 0 and (
            nbytes > getattr(self, 'oneshot_max_bytes', ONESHOT_MAX_BYTES) or sgd is not None or algo is not None)
        if large_ok and sym is not None and sym[1] % 16 == 0 and flat.numel() % 4 == 0:
            buf, off = sym
            use_nvls = (algo or ("nvls" if buf.mc else "p2p")) in ("nvls", "nvls_excl")
            w_arr, w_mc, lr = None, None, 0.0
            if sgd is not None:
                wsym = self.lookup(sgd[0])
                if wsym is None or sgd[0].numel() != flat.numel():
                    raise RuntimeError("fused SGD needs the weights in symmetric memory with the gradient's shape")
                w_arr, w_mc, lr = wsym[0].ptr_array(wsym[1]), (wsym[0].mc + wsym[1]) if wsym[0].mc else 0, float(sgd[1])
                use_nvls = use_nvls and bool(w_mc)
            rc = lib.b200_ar_large_f32(buf.ptr_array(off), (buf.mc + off) if buf.mc else 0, self._flag_arr[channel],
                                       w_arr, w_mc, self.rank, self.size, flat.numel(), scale, lr,
                                       (3 if algo == "nvls_excl" else 2) if use_nvls else 1, blocks or 64,
                                       self._next_epoch(channel), self._stream())
            _native.check(rc, "all_reduce(large)")
        elif nbytes <= min(PUSH_SLOT_BYTES, getattr(self, "oneshot_max_bytes", ONESHOT_MAX_BYTES)) and sgd is None:
            # latency path: one kernel pushes this rank's data into every peer's ring slot, synchronises once and
            # reduces locally - no staging copy, the tensor can live anywhere
            buf = self._push_calls & 1
            self._push_calls += 1
            rc = lib.b200_ar_push(self._push_arr, self._flag_arr[channel], self.rank, self.size,
                                  c_void_p(flat.data_ptr()), c_void_p(flat.data_ptr()), flat.numel(), _DT[t.dtype], op,
                                  scale, PUSH_SLOT_BYTES, buf, max(1, min(16, (nbytes + 16383) >> 14)),
                                  self._next_epoch(channel), self._stream())
            _native.check(rc, "all_reduce(push)")
            if flat.data_ptr() != t.data_ptr():
                t.copy_(flat.view_as(t))
        else:
            if sgd is not None:
                raise RuntimeError("fused SGD is only available on symmetric fp32 buckets")
            # staged path: chunks through the symmetric staging buffer
            st = self._staging
            chunk = (STAGING_BYTES if large_ok else min(STAGING_BYTES, ONESHOT_MAX_BYTES * 16)) // es
            for s in range(0, flat.numel(), chunk):
                part = flat[s:s + chunk]
                n = part.numel()
                stage = st.tensor(n, t.dtype)
                stage.copy_(part)
                if large_ok and n % 4 == 0 and n * es > ONESHOT_MAX_BYTES:
                    use_nvls = (algo or ("nvls" if st.mc else "p2p")) == "nvls"
                    rc = lib.b200_ar_large_f32(st.ptr_array(0), st.mc or 0, self._flag_arr[channel], None, None,
                                               self.rank, self.size, n, scale, 0.0, 2 if use_nvls else 1,
                                               blocks or 64, self._next_epoch(channel), self._stream())
                    _native.check(rc, "all_reduce(staged large)")
                    part.copy_(stage)
                else:
                    rc = lib.b200_ar_oneshot(st.ptr_array(0), self._flag_arr[channel], self.rank, self.size,
                                             c_void_p(part.data_ptr()), n, _DT[t.dtype], op, scale,
                                             self._blocks(n * es), self._next_epoch(channel), self._stream())
                    _native.check(rc, "all_reduce(one-shot)")
            if flat.data_ptr() != t.data_ptr():
                t.copy_(flat.view_as(t))
        from ..distributed import Work
        ev = torch.cuda.Event()
        ev.record()
        return Work(ev)

    @_on_device
    def all_reduce_tiles(self, grad: torch.Tensor, weights: torch.Tensor, ready: torch.Tensor, rows: int,
                         slab_floats: int, target: int, scale: float, lr: float, channel: int = 1,
                         blocks: int = 8, use_nvls: Optional[bool] = None):
        """Tile-driven fused bucket all-reduce + SGD (csrc/allreduce.cu: allreduce_tile_f32_kernel).  ``grad`` [rows, F],
        ``weights`` (same shape) and the slab counters ``ready`` are symmetric tensors; the kernel is meant to be
        launched on a side stream WHILE the producer of ``grad`` (ops.fc_bwd_tiles) runs on the main stream."""
        g, w, r = self.lookup(grad), self.lookup(weights), self.lookup(ready)
        if g is None or w is None or r is None:
            raise RuntimeError("all_reduce_tiles needs gradient, weights and counters in symmetric memory")
        lib = _lib()
        PP = ctypes.POINTER(c_void_p)
        lib.b200_ar_tile_f32.argtypes = [PP, c_void_p, PP, PP, c_void_p, PP, c_int, c_int, c_int, c_size_t, c_size_t,
                                         c_uint32, c_float, c_float, c_int, c_uint32, c_void_p]
        nvls = bool(g[0].mc and w[0].mc) if use_nvls is None else (use_nvls and bool(g[0].mc and w[0].mc))
        F = grad.numel() // rows
        rc = lib.b200_ar_tile_f32(g[0].ptr_array(g[1]), (g[0].mc + g[1]) if nvls else 0, self._flag_arr[channel],
                                  w[0].ptr_array(w[1]), (w[0].mc + w[1]) if nvls else 0, r[0].ptr_array(r[1]),
                                  self.rank, self.size, rows, F, slab_floats, target & 0xFFFFFFFF, scale, lr, blocks,
                                  self._next_epoch(channel), self._stream())
        _native.check(rc, "all_reduce(tiles)")

    @_on_device
    def all_reduce_small_symmetric(self, t: torch.Tensor, scale: float, out: Optional[torch.Tensor] = None,
                                   sgd: Optional[Tuple[torch.Tensor, float]] = None, channel: int = 0,
                                   bcast: Optional[Tuple[torch.Tensor, int]] = None):
        """One-shot fp32 all-reduce reading a SYMMETRIC tensor on every rank directly (no staging copy):
        ``out = scale * sum`` and/or ``sgd[0] -= sgd[1] * scale * sum``; ``t`` itself is left untouched."""
        n = t.numel()
        if self.size == 1:
            if out is not None:
                out.copy_(t * scale)
            if sgd is not None:
                sgd[0].view(-1)[:n].add_(t.view(-1), alpha=-float(sgd[1]) * scale)
            return
        self._ensure(t)
        sym = self.lookup(t)
        assert sym is not None and sym[1] % 16 == 0 and n % 4 == 0 and t.dtype == torch.float32
        buf, off = sym
        lib = _lib()
        args = (buf.ptr_array(off), self._flag_arr[channel], self.rank, self.size,
                c_void_p(out.data_ptr() if out is not None else 0),
                c_void_p(sgd[0].data_ptr() if sgd is not None else 0), n, scale,
                float(sgd[1]) if sgd is not None else 0.0,
                max(1, min(32, (n // 4 + 511) // 512)), self._next_epoch(channel), self._stream())
        if bcast is None:
            rc = lib.b200_ar_oneshot_f32x4(*args)
        else:
            # ``bcast=(symmetric fp32 tensor, root)``: every rank also takes the root's copy of that tensor, in the
            # same kernel (DDP's per-forward BatchNorm-buffer sync folded into the small-bucket all-reduce)
            bt, root = bcast
            bs = self.lookup(bt)
            assert bs is not None and bs[1] % 16 == 0 and bt.dtype == torch.float32 and bt.numel() % 4 == 0
            PP = ctypes.POINTER(c_void_p)
            lib.b200_ar_oneshot_f32x4_bcast.argtypes = [PP, PP, c_int, c_int, c_void_p, c_void_p, c_size_t, c_float,
                                                        c_float, c_int, c_uint32, c_void_p, PP, c_size_t, c_int]
            rc = lib.b200_ar_oneshot_f32x4_bcast(*args, bs[0].ptr_array(bs[1]), bt.numel(), int(root))
        _native.check(rc, "all_reduce(one-shot f32x4)")

    @_on_device
    def broadcast_symmetric_inplace(self, t: torch.Tensor, src: int, channel: int = 0):
        """Every rank overwrites its copy of the symmetric tensor ``t`` with rank ``src``'s (one tiny kernel)."""
        if self.size == 1:
            return
        sym = self.lookup(t)
        nb = t.numel() * t.element_size()
        assert sym is not None and sym[1] % 16 == 0 and nb % 16 == 0
        buf, off = sym
        rc = _lib().b200_ar_broadcast(buf.ptr_array(off), self._flag_arr[channel], self.rank, self.size,
                                      c_void_p(t.data_ptr()), nb, src, self._blocks(nb), self._next_epoch(channel),
                                      self._stream())
        _native.check(rc, "broadcast(symmetric)")

    def _padded_io(self, t: torch.Tensor):
        """(contiguous byte view padded to 16 B, needs_copy_back)"""
        flat = t.contiguous().view(-1).view(torch.uint8)
        if flat.numel() % 16 == 0 and flat.data_ptr() % 16 == 0 and t.is_contiguous():
            return flat, False
        pad = torch.zeros((flat.numel() + 15) // 16 * 16, dtype=torch.uint8, device=t.device)
        pad[:flat.numel()].copy_(flat)
        return pad, True

    @_on_device
    def broadcast(self, t: torch.Tensor, src: int, channel: int = 0):
        if self.size == 1:
            return None
        if not t.is_cuda:
            raise RuntimeError("broadcast on the native backend needs a CUDA tensor (use backend='gloo' for CPU)")
        self._ensure(t)
        io, copy_back = self._padded_io(t)
        lib, st = _lib(), self._staging
        for s in range(0, io.numel(), STAGING_BYTES):
            part = io[s:s + STAGING_BYTES]
            if self.rank == src:
                st.bytes_view[:part.numel()].copy_(part)
            rc = lib.b200_ar_broadcast(st.ptr_array(0), self._flag_arr[channel], self.rank, self.size,
                                       c_void_p(part.data_ptr()), part.numel(), src, self._blocks(part.numel()),
                                       self._next_epoch(channel), self._stream())
            _native.check(rc, "broadcast")
        if copy_back:
            t.copy_(io[:t.numel() * t.element_size()].view(t.dtype).view(t.shape))
        return None

    @_on_device
    def all_gather(self, out: torch.Tensor, t: torch.Tensor, channel: int = 0):
        if self.size == 1:
            out.view(-1).copy_(t.view(-1))
            return None
        if not t.is_cuda:
            raise RuntimeError("all_gather on the native backend needs CUDA tensors (use backend='gloo' for CPU)")
        self._ensure(t)
        nb = t.numel() * t.element_size()
        if nb > STAGING_BYTES:
            raise RuntimeError("all_gather: per-rank tensor exceeds the 32 MiB staging buffer")
        pad = (nb + 15) // 16 * 16
        st = self._staging
        st.bytes_view[:nb].copy_(t.contiguous().view(-1).view(torch.uint8))
        tmp = torch.empty(self.size * pad, dtype=torch.uint8, device=t.device)
        rc = _lib().b200_ar_allgather(st.ptr_array(0), self._flag_arr[channel], self.rank, self.size,
                                      c_void_p(tmp.data_ptr()), pad, self._blocks(pad), self._next_epoch(channel),
                                      self._stream())
        _native.check(rc, "all_gather")
        out.view(-1).view(torch.uint8).copy_(tmp.view(self.size, pad)[:, :nb].reshape(-1))
        return None

    def barrier(self, channel: int = 0):
        self._ensure()
        if self.size == 1:
            torch.cuda.synchronize(self.device)
            return
        with torch.cuda.device(self.device):
            rc = _lib().b200_ar_barrier(self._flag_arr[channel], self.rank, self.size, self._next_epoch(channel),
                                        self._stream())
            _native.check(rc, "barrier")
            torch.cuda.current_stream().synchronize()

    def close(self):
        if self._comm:
            try:
                torch.cuda.synchronize(self.device)
            except Exception:
                pass
            for b in reversed(self._buffers):
                b.free()
            self._buffers = []
            _lib().b200_symm_comm_destroy(self._comm)
            self._comm = None
