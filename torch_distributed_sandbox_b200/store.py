"""Python face of the native TCP key-value store (csrc/store.cpp).

Plays the role torch's C++ ``TCPStore`` plays under the reference's
``dist.init_process_group`` calls (test_init.py:91, allreduce_toy.py:44,
mnist_distributed.py:50): rank 0 hosts it on MASTER_ADDR:MASTER_PORT, every rank connects.
"""
from __future__ import annotations

import ctypes
from ctypes import c_char_p, c_int, c_longlong, c_void_p
from typing import Optional

from . import _native


def _lib():
    lib = _native.host()
    if not getattr(lib, "_b200_store_bound", False):
        lib.b200h_last_error.restype = c_char_p
        lib.b200h_store_server_start.restype = c_void_p
        lib.b200h_store_server_start.argtypes = [c_char_p, c_int, ctypes.POINTER(c_int)]
        lib.b200h_store_server_stop.argtypes = [c_void_p]
        lib.b200h_store_connect.restype = c_void_p
        lib.b200h_store_connect.argtypes = [c_char_p, c_int, c_int]
        lib.b200h_store_close.argtypes = [c_void_p]
        lib.b200h_store_set.argtypes = [c_void_p, c_char_p, c_char_p, c_int]
        lib.b200h_store_get.argtypes = [c_void_p, c_char_p, c_void_p, c_int, ctypes.POINTER(c_int), c_int]
        lib.b200h_store_add.argtypes = [c_void_p, c_char_p, c_longlong, ctypes.POINTER(c_longlong)]
        lib.b200h_store_check.argtypes = [c_void_p, c_char_p]
        lib.b200h_store_delete.argtypes = [c_void_p, c_char_p]
        lib.b200h_store_num_keys.argtypes = [c_void_p]
        lib.b200h_store_num_keys.restype = c_longlong
        lib._b200_store_bound = True
    return lib


class StoreError(RuntimeError):
    pass


class StoreTimeout(StoreError, TimeoutError):
    pass


def _err() -> str:
    m = _lib().b200h_last_error()
    return m.decode() if m else "?"


class TCPStore:
    """``TCPStore(host, port, is_master)``; ``port=0`` on the master picks a free port
    (read it back from ``.port``).  Values are ``bytes``; ``get`` blocks up to ``timeout``."""

    def __init__(self, host: str, port: int, is_master: bool = False, timeout: float = 300.0,
                 prefix: str = ""):
        lib = _lib()
        self.host, self.timeout, self.prefix = host, float(timeout), prefix
        self._server = None
        self._client = None
        if is_master:
            actual = c_int(0)
            self._server = lib.b200h_store_server_start(host.encode(), int(port), ctypes.byref(actual))
            if not self._server:
                raise StoreError(_err())
            port = actual.value
        self.port = int(port)
        self._client = lib.b200h_store_connect(host.encode(), self.port, int(self.timeout * 1000))
        if not self._client:
            msg = _err()
            self.close()
            raise StoreTimeout(msg)

    # -- key/value -------------------------------------------------------------
    def _k(self, key: str) -> bytes:
        return (self.prefix + key).encode()

    def set(self, key: str, value) -> None:
        if isinstance(value, str):
            value = value.encode()
        if _lib().b200h_store_set(self._client, self._k(key), bytes(value), len(value)) != 0:
            raise StoreError(_err())

    def get(self, key: str, timeout: Optional[float] = None) -> bytes:
        t = self.timeout if timeout is None else timeout
        cap = 1 << 16
        while True:
            buf = ctypes.create_string_buffer(cap)
            n = c_int(0)
            rc = _lib().b200h_store_get(self._client, self._k(key), buf, cap, ctypes.byref(n), int(t * 1000))
            if rc == 0:
                return buf.raw[:n.value]
            if rc == 3:                     # buffer too small: retry with the reported size
                cap = n.value
                continue
            if rc == 1:
                raise StoreTimeout(_err())
            raise StoreError(_err())

    def add(self, key: str, delta: int) -> int:
        out = c_longlong(0)
        if _lib().b200h_store_add(self._client, self._k(key), int(delta), ctypes.byref(out)) != 0:
            raise StoreError(_err())
        return int(out.value)

    def check(self, key: str) -> bool:
        return _lib().b200h_store_check(self._client, self._k(key)) == 1

    def delete_key(self, key: str) -> bool:
        return _lib().b200h_store_delete(self._client, self._k(key)) == 1

    def num_keys(self) -> int:
        return int(_lib().b200h_store_num_keys(self._client))

    def wait(self, keys, timeout: Optional[float] = None) -> None:
        for k in keys:
            self.get(k, timeout)

    # -- derived primitives ------------------------------------------------------
    def barrier(self, name: str, world_size: int, timeout: Optional[float] = None) -> None:
        """Store-side barrier (cold path: bring-up / teardown)."""
        n = self.add(f"{name}/count", 1)
        if n == world_size:
            self.set(f"{name}/done", b"1")
        self.get(f"{name}/done", timeout)

    def close(self) -> None:
        lib = _lib()
        if self._client:
            lib.b200h_store_close(self._client)
            self._client = None
        if self._server:
            lib.b200h_store_server_stop(self._server)
            self._server = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass
