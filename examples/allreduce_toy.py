#!/usr/bin/env python
"""All-reduce toy - the counterpart of the reference's ``allreduce_toy.py``: every rank draws
a random integer per step, the ranks sum them with ``all_reduce`` and meet at a barrier.

Same flags as the reference (allreduce_toy.py:60-71); the ones it parses and then ignores
(``--backend``, ``--init-method``, ``--rank``, ``--steps``) work here:

    python examples/allreduce_toy.py -s 2 --backend gloo            # CPU plumbing config
    python examples/allreduce_toy.py -s 8                           # one rank per GPU, our NVLink kernels
    python examples/allreduce_toy.py -s 2 -r 0 -i tcp://10.0.0.1:23456 --backend gloo   # one rank of a manual launch
"""
import argparse
import os
import sys
from random import randint

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import torch_distributed_sandbox_b200 as b2
from torch_distributed_sandbox_b200 import distributed as dist


def run(world_size, rank, steps):
    on_gpu = dist.get_backend() == "native"
    device = torch.device("cuda", rank % max(torch.cuda.device_count(), 1)) if on_gpu else torch.device("cpu")
    for step in range(1, steps + 1):
        value = randint(0, 10)
        # the reference builds a new group every step (allreduce_toy.py:26-27); here the same rank
        # tuple maps to one cached group, so this is a dictionary lookup, not a new communicator
        group = dist.new_group(ranks=list(range(world_size)))
        tensor = torch.tensor(value, dtype=torch.int).to(device)
        dist.all_reduce(tensor, op=dist.ReduceOp.SUM, group=group)
        dist.barrier()
        if rank in (0, 1):
            print("rank: {}, step: {}, value: {}, reduced sum: {}.".format(rank, step, value, tensor.item()))


def setup(rank, world_size, backend="nccl", steps=10, init_method=None):
    if rank == -1:   # serial code
        return
    # as in the reference (allreduce_toy.py:41-48) no torch.cuda.set_device: the group binds to the device of the
    # first collective's tensor (cuda:rank), like ProcessGroupNCCL
    dist.init_process_group(backend, init_method=init_method, rank=rank, world_size=world_size)
    print(f"--> done setting up rank={rank}")
    run(world_size, rank, steps)
    dist.destroy_process_group()


def main(argv=None):
    parser = argparse.ArgumentParser()
    parser.add_argument("--backend", type=str, default="nccl", help="nccl (our NVLink kernels) or gloo (CPU).")
    parser.add_argument("-i", "--init-method", type=str, default=None,
                        help="URL specifying how to initialize the package (default: env:// on a free local port).")
    parser.add_argument("-s", "--world_size", type=int, required=True, help="Number of processes participating in the job.")
    parser.add_argument("-r", "--rank", type=int, default=None,
                        help="Run only this rank in this process (manual launch); default: spawn all ranks.")
    parser.add_argument("--steps", type=int, default=10)
    args = parser.parse_args(argv)
    backend = args.backend
    if backend == "nccl" and not torch.cuda.is_available():
        print("no CUDA device: falling back to the gloo (host) backend")
        backend = "gloo"
    if args.init_method is None:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if args.rank is None:
            os.environ["MASTER_PORT"] = b2.find_free_port()
    if args.rank is not None:
        setup(args.rank, args.world_size, backend, args.steps, args.init_method)
    else:
        b2.spawn(setup, args=(args.world_size, backend, args.steps, args.init_method), nprocs=args.world_size)


if __name__ == "__main__":
    main()
