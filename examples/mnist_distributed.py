#!/usr/bin/env python
"""Data-parallel MNIST training at 3000x3000 - the counterpart of the reference's
``mnist_distributed.py``: one process per GPU, batch 5 per GPU (effective batch 5 x world,
the README's "OOM workaround"), DistributedSampler sharding, DDP gradient averaging.

Same flags (-n/--nodes, -g/--gpus, -nr/--nr, --epochs) and log lines.  The gradient
all-reduce is our NVLink/NVSwitch kernel overlapped with backward; no NCCL call is made.
``--backend gloo --model eager`` runs the plain-PyTorch model on CPU through the same DDP
wrapper (plumbing configuration, tiny images).
"""
import argparse
import os
import sys
from datetime import datetime

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch import nn

import torch_distributed_sandbox_b200 as b2
from torch_distributed_sandbox_b200 import distributed as dist
from torch_distributed_sandbox_b200.data.mnist import DataLoader, open_dataset
from torch_distributed_sandbox_b200.data.resize import resize_u8_reference
from torch_distributed_sandbox_b200.parallel import DistributedDataParallel, DistributedSampler

IMAGE_SHAPE = [3000, 3000]


def train(gpu, args):
    rank = args.nr * args.gpus + gpu
    on_gpu = args.backend == "nccl"
    dist.init_process_group(backend=args.backend, world_size=args.world_size, rank=rank)   # mnist_distributed.py:50
    torch.manual_seed(0)
    if on_gpu:
        torch.cuda.set_device(gpu)                  # after init, as in the reference (:57): the group binds lazily
    batch_size = args.batch_size                    # per GPU: effective batch = batch_size * world_size
    if args.model == "fused":
        from torch_distributed_sandbox_b200.models.module import FusedModule
        model = FusedModule(args.image_shape, batch_size=batch_size, device=torch.device("cuda", gpu), lr=1e-4, seed=0,
                            precision=args.precision)
        model = DistributedDataParallel(model, device_ids=[gpu], gradient_dtype=args.grad_dtype)
        criterion = optimizer = None
    else:
        from torch_distributed_sandbox_b200.models import reference_init
        model = reference_init(args.image_shape, seed=0)
        if on_gpu:
            model.cuda(gpu)
        criterion = nn.CrossEntropyLoss()
        optimizer = torch.optim.SGD(model.parameters(), 1e-4)
        model = DistributedDataParallel(model, device_ids=[gpu] if on_gpu else None)
    dataset = open_dataset(args.data, synthetic=args.synthetic, quiet=(gpu != 0))
    sampler = DistributedSampler(dataset, num_replicas=args.world_size, rank=rank)
    loader = DataLoader(dataset, batch_size=batch_size, shuffle=False, pin_memory=on_gpu, sampler=sampler)
    start = datetime.now()
    total_step = len(loader) if args.max_steps is None else min(len(loader), args.max_steps)
    for epoch in range(args.epochs):
        for i, (images, labels) in enumerate(loader):
            if i >= total_step:
                break
            if args.model == "fused":
                loss = model.train_step(images.cuda(non_blocking=True), labels.cuda(non_blocking=True))
            else:
                x = torch.from_numpy(resize_u8_reference(images.numpy(), args.image_shape)).float().div(255)[:, None]
                if on_gpu:
                    x, labels = x.cuda(non_blocking=True), labels.cuda(non_blocking=True)
                loss = criterion(model(x), labels)
                optimizer.zero_grad()
                loss.backward()
                optimizer.step()
            # the reference also builds an (unused) group here every step (mnist_distributed.py:99-100)
            dist.new_group(ranks=list(range(args.world_size)))
            if (i + 1) % args.log_every == 0 and gpu == 0:
                print("Rank [{}], Epoch [{}/{}], Step [{}/{}], Loss: {:.4f}".format(
                    rank, epoch + 1, args.epochs, i + 1, total_step, loss.item()))
    if on_gpu:
        torch.cuda.synchronize()
    if gpu == 0:
        print("Training complete in: " + str(datetime.now() - start))
        if args.save:
            torch.save(model.state_dict(), args.save)     # keys carry DDP's "module." prefix
    dist.destroy_process_group()


def parse(argv=None):
    parser = argparse.ArgumentParser()
    parser.add_argument("-n", "--nodes", default=1, type=int, metavar="N", help="number of nodes (default: 1)")
    parser.add_argument("-g", "--gpus", default=1, type=int, help="number of gpus per node")
    parser.add_argument("-nr", "--nr", default=0, type=int, help="ranking within the nodes")
    parser.add_argument("--epochs", default=2, type=int, metavar="N", help="number of total epochs to run")
    parser.add_argument("--image-shape", default=IMAGE_SHAPE, type=int, nargs=2)
    parser.add_argument("--batch-size", default=5, type=int)
    parser.add_argument("--backend", default="nccl", choices=["nccl", "gloo"])
    parser.add_argument("--model", default=None, choices=["fused", "eager"])
    parser.add_argument("--data", default="./data")
    parser.add_argument("--synthetic", default=None, type=int,
                        help="train on N generated digits (default: the idx files under --data when present, "
                             "else 60000 generated digits)")
    parser.add_argument("--max-steps", default=None, type=int)
    parser.add_argument("--log-every", default=100, type=int)
    parser.add_argument("--grad-dtype", default="fp32", choices=["fp32", "bf16"],
                        help="fused model: dtype of the fc.weight gradient exchange (bf16 = half the all-reduce bytes, "
                             "fp32 accumulation in the switch; like torch DDP's bf16_compress_hook)")
    parser.add_argument("--precision", default="tf32", choices=["tf32", "bf16"],
                        help="fused model: tf32 = fp32 storage + TF32 conv math (the reference's numerics, default); "
                             "bf16 = bf16 activations/gradients (faster, lower precision)")
    parser.add_argument("--save", default=None)
    args = parser.parse_args(argv)
    args.world_size = args.gpus * args.nodes
    if args.backend == "nccl" and not torch.cuda.is_available():
        print("no CUDA device: using --backend gloo --model eager")
        args.backend = "gloo"
    if args.model is None:
        args.model = "fused" if args.backend == "nccl" else "eager"
    return args


if __name__ == "__main__":
    args = parse()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ["MASTER_PORT"] = b2.find_free_port()
    b2.spawn(train, nprocs=args.gpus, args=(args,))
