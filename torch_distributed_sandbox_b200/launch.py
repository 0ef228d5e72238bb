"""Process launcher (L1 / D1 in SURVEY.md): one OS process per rank.

Reference behaviour being provided: ``mp.spawn(fn, args=..., nprocs=N)`` as used at
test_init.py:116, allreduce_toy.py:74 and mnist_distributed.py:127 - ``fn(i, *args)`` for
``i in range(nprocs)``, start method *spawn*, the parent's environment inherited by the
children (the scripts export MASTER_ADDR/MASTER_PORT before spawning), ``join=True``, and on
the first failure the surviving siblings are torn down and the child's traceback is
re-raised in the parent.  Also the small helpers of test_init.py:18-53.

This is our own implementation on the stdlib ``multiprocessing`` (no torch.multiprocessing):
children report through a pipe, the parent polls sentinels, SIGTERM then SIGKILL.
"""
from __future__ import annotations

import multiprocessing as _mp
import os
import signal
import socket
import sys
import time
import traceback
from contextlib import closing
from typing import Callable, Optional, Sequence


class ProcessRaisedError(RuntimeError):
    """A spawned rank raised; carries the rank index and the child's formatted traceback."""

    def __init__(self, msg: str, error_index: int, error_pid: int):
        super().__init__(msg)
        self.error_index = error_index
        self.error_pid = error_pid


class ProcessExitedError(RuntimeError):
    """A spawned rank died without reporting an exception (signal or sys.exit(!=0))."""

    def __init__(self, msg: str, error_index: int, error_pid: int, exit_code: int):
        super().__init__(msg)
        self.error_index = error_index
        self.error_pid = error_pid
        self.exit_code = exit_code


def find_free_port() -> str:
    """A TCP port that was free a moment ago, as ``str`` (the reference returns a string and
    feeds it to ``os.environ`` directly: test_init.py:45-53).  SO_REUSEADDR is set *before*
    bind here; the unavoidable close-then-rebind window is narrowed by the store retrying."""
    with closing(socket.socket(socket.AF_INET, socket.SOCK_STREAM)) as s:
        s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        s.bind(("127.0.0.1", 0))
        return str(s.getsockname()[1])


def set_sharing_strategy(new_strategy: Optional[str] = None) -> str:
    """Tensor-sharing strategy for worker processes (test_init.py:18-34): the given one, else
    ``file_system`` on macOS and ``file_descriptor`` elsewhere.  Returns the strategy set."""
    import torch.multiprocessing as tmp
    if new_strategy is None:
        new_strategy = "file_system" if sys.platform == "darwin" else "file_descriptor"
    tmp.set_sharing_strategy(new_strategy)
    return new_strategy


def use_file_system_sharing_strategy() -> str:
    """Force ``file_system`` sharing - the "too many open file descriptors" escape hatch
    (test_init.py:36-43)."""
    return set_sharing_strategy("file_system")


def _child_main(fn, index, args, err_conn):
    try:
        fn(index, *args)
    except KeyboardInterrupt:
        pass
    except BaseException:
        try:
            err_conn.send(traceback.format_exc())
        finally:
            err_conn.close()
        sys.exit(1)
    finally:
        sys.stdout.flush()
        sys.stderr.flush()


class SpawnContext:
    def __init__(self, procs, conns):
        self.processes = procs
        self._conns = conns

    def pids(self):
        return [p.pid for p in self.processes]

    def _kill_survivors(self, grace: float = 5.0) -> None:
        alive = [p for p in self.processes if p.is_alive()]
        for p in alive:
            p.terminate()
        deadline = time.time() + grace
        for p in alive:
            p.join(max(0.0, deadline - time.time()))
        for p in alive:
            if p.is_alive():
                os.kill(p.pid, signal.SIGKILL)
                p.join()

    def join(self, timeout: Optional[float] = None) -> bool:
        """True when every rank exited cleanly; raises on the first failure."""
        deadline = None if timeout is None else time.time() + timeout
        while True:
            failed = None
            running = False
            for i, p in enumerate(self.processes):
                if p.exitcode is None:
                    running = True
                elif p.exitcode != 0 and failed is None:
                    failed = i
            if failed is not None:
                p = self.processes[failed]
                tb = None
                if self._conns[failed].poll(0.2):
                    try:
                        tb = self._conns[failed].recv()
                    except EOFError:
                        tb = None
                self._kill_survivors()
                if tb is not None:
                    raise ProcessRaisedError(f"\n\n-- Process {failed} terminated with the following error:\n{tb}",
                                             failed, p.pid)
                if p.exitcode < 0:
                    name = signal.Signals(-p.exitcode).name
                    raise ProcessExitedError(f"process {failed} terminated with signal {name}", failed, p.pid,
                                             p.exitcode)
                raise ProcessExitedError(f"process {failed} terminated with exit code {p.exitcode}", failed, p.pid,
                                         p.exitcode)
            if not running:
                return True
            if deadline is not None and time.time() > deadline:
                return False
            _mp.connection.wait([p.sentinel for p in self.processes if p.exitcode is None], timeout=0.25)


def spawn(fn: Callable, args: Sequence = (), nprocs: int = 1, join: bool = True, daemon: bool = False,
          start_method: str = "spawn"):
    """Run ``fn(i, *args)`` in ``nprocs`` fresh processes (i = 0..nprocs-1)."""
    if nprocs < 1:
        raise ValueError("nprocs must be >= 1")
    ctx = _mp.get_context(start_method)
    procs, conns = [], []
    for i in range(nprocs):
        parent_conn, child_conn = ctx.Pipe(duplex=False)
        p = ctx.Process(target=_child_main, args=(fn, i, tuple(args), child_conn), daemon=daemon)
        p.start()
        child_conn.close()
        procs.append(p)
        conns.append(parent_conn)
    sc = SpawnContext(procs, conns)
    if not join:
        return sc
    while not sc.join():
        pass
    return None
