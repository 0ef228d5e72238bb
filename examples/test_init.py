#!/usr/bin/env python
"""Process-group bring-up smoke test - the counterpart of the reference's ``test_init.py``
(test_init.py:55-121): spawn 4 processes, rendezvous over env:// on 127.0.0.1:<free port>,
return from ``init_process_group``, exit.  Differences that are deliberate: the group is
torn down with ``cleanup`` (the reference defines it but never calls it), and with GPUs
present a collective is actually run so the NVLink path - not just the TCP store - is proven.

    python examples/test_init.py [--world_size 4] [--backend gloo|nccl]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import torch_distributed_sandbox_b200 as b2
from torch_distributed_sandbox_b200 import distributed as dist
from torch_distributed_sandbox_b200.data.mnist import get_batch  # noqa: F401  (re-exported helper)
from torch_distributed_sandbox_b200.launch import (find_free_port, set_sharing_strategy,  # noqa: F401
                                                   use_file_system_sharing_strategy)


def setup_process(rank, world_size, port, backend="gloo"):
    """Initialise the distributed environment of one rank; ``rank == -1`` means serial code."""
    if rank == -1:
        return
    print(f"setting up rank={rank} (with world_size={world_size})")
    master_addr = "127.0.0.1"
    os.environ["MASTER_ADDR"] = master_addr
    print(f"MASTER_ADDR={master_addr!r}")
    os.environ["MASTER_PORT"] = port
    print(f"port={port!r}")
    if torch.cuda.is_available() and torch.cuda.device_count() >= world_size:
        backend = "nccl"
        torch.cuda.set_device(rank)
    print(f"backend={backend!r}")
    dist.init_process_group(backend, rank=rank, world_size=world_size)
    print(f"--> done setting up rank={rank}")


def cleanup(rank):
    """Destroy the process group unless the code runs serially."""
    if rank != -1:
        dist.destroy_process_group()


def _worker(rank, world_size, port, backend):
    setup_process(rank, world_size, port, backend)
    t = torch.ones(1, device="cuda" if dist.get_backend() == "native" else "cpu")
    dist.all_reduce(t)
    assert int(t.item()) == world_size
    cleanup(rank)


def test_setup(world_size=4, backend="gloo"):
    print("test_setup")
    port = find_free_port()
    b2.spawn(_worker, args=(world_size, port, backend), nprocs=world_size)
    print("successful test_setup!")


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--world_size", type=int, default=4)
    ap.add_argument("--backend", default="gloo")
    a = ap.parse_args()
    test_setup(a.world_size, a.backend)
