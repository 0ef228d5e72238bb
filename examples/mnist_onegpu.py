#!/usr/bin/env python
"""Single-GPU MNIST training at 3000x3000 - the counterpart of the reference's
``mnist_onegpu.py`` (same CNN, batch 5, SGD lr 1e-4, same log lines, ``--epochs`` flag).

What changed under the hood: digits stay uint8 28x28 until they are on the device, the
resize is a CUDA kernel inside the step, and the step is the fused sm_100a engine.  There is
no network here, so ``--synthetic N`` (default 60000) generates MNIST-shaped data; point
``--data`` at a folder holding ``MNIST/raw/*-ubyte`` to use real files.
"""
import argparse
import os
import sys
from datetime import datetime

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from torch_distributed_sandbox_b200.data.mnist import DataLoader, open_dataset

IMAGE_SHAPE = [3000, 3000]


def train(gpu, args):
    from torch_distributed_sandbox_b200.models.module import FusedModule
    torch.manual_seed(0)
    torch.cuda.set_device(gpu)
    batch_size = args.batch_size            # 5 in the reference (10 OOMs a 24 GB card: mnist_onegpu.py:44-45)
    model = FusedModule(args.image_shape, batch_size=batch_size, device=torch.device("cuda", gpu), lr=1e-4, seed=0,
                            precision=args.precision)
    dataset = open_dataset(args.data, synthetic=args.synthetic, quiet=(gpu != 0))
    loader = DataLoader(dataset, batch_size=batch_size, shuffle=True, pin_memory=True)
    start = datetime.now()
    total_step = len(loader) if args.max_steps is None else min(len(loader), args.max_steps)
    for epoch in range(args.epochs):
        for i, (images, labels) in enumerate(loader):
            if i >= total_step:
                break
            images = images.cuda(non_blocking=True)
            labels = labels.cuda(non_blocking=True)
            loss = model.train_step(images, labels)      # forward + loss + backward + SGD, fused
            if (i + 1) % args.log_every == 0 and gpu == 0:
                print("Epoch [{}/{}], Step [{}/{}], Loss: {:.4f}".format(epoch + 1, args.epochs, i + 1, total_step,
                                                                         loss.item()))
    torch.cuda.synchronize()
    if gpu == 0:
        print("Training complete in: " + str(datetime.now() - start))
    if args.save:
        torch.save(model.state_dict(), args.save)         # loadable by the reference ConvNet


def parse(argv=None):
    parser = argparse.ArgumentParser()
    parser.add_argument("--epochs", default=2, type=int, metavar="N", help="number of total epochs to run")
    parser.add_argument("--image-shape", default=IMAGE_SHAPE, type=int, nargs=2)
    parser.add_argument("--batch-size", default=5, type=int)
    parser.add_argument("--data", default="./data")
    parser.add_argument("--synthetic", default=None, type=int,
                        help="train on N generated digits (default: the idx files under --data when present, "
                             "else 60000 generated digits)")
    parser.add_argument("--max-steps", default=None, type=int)
    parser.add_argument("--log-every", default=100, type=int)
    parser.add_argument("--precision", default="tf32", choices=["tf32", "bf16"],
                        help="fused model: tf32 = fp32 storage + TF32 conv math (the reference's numerics, default); "
                             "bf16 = bf16 activations/gradients (faster, lower precision)")
    parser.add_argument("--save", default=None, help="write a torch state_dict checkpoint here")
    args = parser.parse_args(argv)
    return args


if __name__ == "__main__":
    train(0, parse())
