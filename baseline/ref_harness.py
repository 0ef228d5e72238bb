"""Reference arm of bench.py: drives the UNMODIFIED reference installed in baseline/_ref.

Three measurements, all with the reference's own classes and call sequence
(mnist_distributed.py:48-109), none of this repo's models / kernels / engine on the path:

  value       device-timed training step: ``ConvNet`` + ``CrossEntropyLoss`` + ``SGD(1e-4)`` +
              ``DistributedDataParallel`` built exactly as ``train()`` builds them, fed
              device-resident synthetic 3000x3000 batches (the GPU-step baseline);
  e2e.value   same step, but every step copies its fp32 batch from pinned host memory
              (``images.cuda(non_blocking=True)``, 180 MB at batch 5) and reads ``loss.item()``
              back - generous to the reference: its CPU PIL resize is NOT charged;
  e2e_stock_loader   the stock ``train(gpu, args)`` function called as is (torchvision MNIST +
              Resize + ToTensor + DataLoader + per-step new_group), on idx files pre-seeded
              with exactly enough synthetic digits for the requested number of steps; the
              time is the reference's own "Training complete in:" print.
"""
from __future__ import annotations

import argparse
import contextlib
import io
import json
import os
import re
import sys
import tempfile


def _seed_idx(root, n):
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, repo)
    from torch_distributed_sandbox_b200.data.mnist import write_synthetic_idx   # data files only
    write_synthetic_idx(root, n_train=n, n_test=10, seed=7)


def _stock_train(ref, local_rank, world, steps, batch):
    """Call the reference's train() unmodified for exactly `steps` iterations."""
    import torch.distributed as dist
    tmp = tempfile.mkdtemp(prefix=f"b200dp_ref_r{local_rank}_")
    _seed_idx(os.path.join(tmp, "data"), steps * batch * world)
    cwd = os.getcwd()
    os.chdir(tmp)
    args = argparse.Namespace(nodes=1, gpus=world, nr=0, epochs=1, world_size=world)
    buf = io.StringIO()
    try:
        with contextlib.redirect_stdout(buf):
            ref.train(local_rank, args)
    finally:
        os.chdir(cwd)
        if dist.is_initialized():
            dist.destroy_process_group()
    m = re.search(r"Training complete in: (\d+):(\d+):([\d.]+)", buf.getvalue())
    if not m:
        return None
    return int(m.group(1)) * 3600 + int(m.group(2)) * 60 + float(m.group(3))


def _stock_main(argv):
    """Child process of the stock-loader stage (world 1 only): its own rendezvous on a fresh port
    for every train() call, nothing inherited from a torchrun agent.  Prints one JSON line."""
    ref_dir, steps, batch, image = argv[0], int(argv[1]), int(argv[2]), int(argv[3])
    for k in list(os.environ):
        if k.startswith(("TORCHELASTIC_", "TORCH_NCCL_", "PET_")) or k in (
                "RANK", "LOCAL_RANK", "WORLD_SIZE", "GROUP_RANK", "ROLE_RANK", "LOCAL_WORLD_SIZE", "ROLE_WORLD_SIZE",
                "GROUP_WORLD_SIZE", "ROLE_NAME", "MASTER_PORT", "MASTER_ADDR"):
            del os.environ[k]
    sys.path.insert(0, ref_dir)
    import mnist_distributed as ref
    if image != 3000:
        ref.IMAGE_SHAPE = [image, image]
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    out = {}
    for tag, n in (("warmup", 2), ("timed", steps)):
        os.environ["MASTER_PORT"] = ref.find_free_port()
        out[tag] = _stock_train(ref, 0, 1, n, batch)
    print(json.dumps(out), flush=True)


def _stock_stage(ref_dir, a, local_rank, timeout_s):
    """Stock train() (torchvision loader + PIL resize) timed in a fresh process under a hard timeout.
    A process group is never re-initialised on the launcher's store: round 1 did that after a
    destroy_process_group() and hung every world > 1 run (ranks dialled a closed NCCL bootstrap socket)."""
    import subprocess
    env = dict(os.environ)
    vis = env.get("CUDA_VISIBLE_DEVICES")
    env["CUDA_VISIBLE_DEVICES"] = vis.split(",")[local_rank] if vis else str(local_rank)
    cmd = [sys.executable, os.path.abspath(__file__), "--stock", ref_dir, str(a.stock_loader_steps), str(a.batch),
           str(a.image)]
    try:
        r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout_s)
    except subprocess.TimeoutExpired:
        return {"error": f"stock train() did not finish within {timeout_s} s (killed)"}
    line = next((ln for ln in reversed(r.stdout.splitlines()) if ln.startswith("{")), None)
    if r.returncode != 0 or line is None:
        return {"error": f"rc={r.returncode}: {(r.stderr or r.stdout)[-300:]}"}
    secs = json.loads(line).get("timed")
    if not secs:
        return {"error": "no 'Training complete in' line"}
    return {"value": a.stock_loader_steps * a.batch / secs, "unit": "img/s", "steps": a.stock_loader_steps,
            "seconds": secs, "world": 1,
            "note": "stock train(): torchvision MNIST + PIL Resize(3000) + ToTensor on the CPU, num_workers=0, "
                    "180 MB H2D per step, per-step new_group; own process, own rendezvous"}


def run(a, ref_dir, ranks, timed_loop):
    rank, local_rank, world = ranks
    sys.path.insert(0, ref_dir)
    import torch
    import torch.distributed as dist
    import torch.nn as nn
    import torch.nn.functional as F
    import mnist_distributed as ref                       # the unmodified reference module

    if world == 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", ref.find_free_port())
    torch.cuda.set_device(local_rank)
    B, S = a.batch, a.image
    if S != 3000:
        ref.IMAGE_SHAPE = [S, S]                          # the reference's own config constant (R12)

    # ---- build exactly as train() does (mnist_distributed.py:50-67) -----------------
    dist.init_process_group(backend="nccl", world_size=world, rank=rank)
    torch.manual_seed(0)
    model = ref.ConvNet()
    model(torch.zeros([1, 1] + ref.IMAGE_SHAPE))
    model.cuda(local_rank)
    criterion = nn.CrossEntropyLoss().cuda(local_rank)
    optimizer = torch.optim.SGD(model.parameters(), 1e-4)
    model = nn.parallel.DistributedDataParallel(model, device_ids=[local_rank])

    nb = 2
    g = torch.Generator().manual_seed(100 + rank)
    small = torch.rand(nb, B, 1, 28, 28, generator=g)
    dev_imgs = torch.stack([F.interpolate(small[i].cuda(), size=(S, S), mode="bilinear") for i in range(nb)])
    dev_labs = torch.randint(0, 10, (nb, B), generator=g).cuda()
    host_imgs = dev_imgs.cpu().pin_memory()
    host_labs = dev_labs.cpu().pin_memory()

    def sync_all():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        t = torch.tensor([v], dtype=torch.float32, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    last = {}

    def one_step(images, labels):                         # mnist_distributed.py:90-97
        outputs = model(images)
        loss = criterion(outputs, labels)
        optimizer.zero_grad()
        loss.backward()
        optimizer.step()
        return loss

    def dev_step(i):
        last["loss"] = one_step(dev_imgs[i % nb], dev_labs[i % nb])

    def e2e_step(i):                                      # mnist_distributed.py:87-88 + loss.item()
        images = host_imgs[i % nb].cuda(non_blocking=True)
        labels = host_labs[i % nb].cuda(non_blocking=True)
        last["loss_val"] = one_step(images, labels).item()

    ms, clocks = timed_loop(dev_step, a.steps, a.warmup, sync_all, max_over_ranks, local_rank)
    ms_e2e, _ = timed_loop(e2e_step, a.steps, max(3, a.warmup // 2), sync_all, max_over_ranks, local_rank,
                           sample_clocks=False)
    peak = torch.cuda.max_memory_allocated() / 2 ** 30
    final_loss = last.get("loss_val")
    if rank == 0:                                         # the measured numbers, on stderr at once: never lost
        print(f"[ref_harness] world={world} ms_per_step={ms:.3f} e2e_ms={ms_e2e:.3f}", file=sys.stderr, flush=True)
    dist.barrier()
    dist.destroy_process_group()                          # last use of the launcher's store in this process
    del model, optimizer, dev_imgs, host_imgs
    torch.cuda.empty_cache()

    stock = None
    if a.stock_loader_steps > 0 and world == 1:
        stock = _stock_stage(ref_dir, a, local_rank, timeout_s=int(os.environ.get("B200DP_STOCK_TIMEOUT", "90")))
    elif a.stock_loader_steps > 0:
        stock = {"skipped": "stock train() is timed at world 1 only (see e2e_stock_loader of the 1-GPU line)"}
    if rank != 0:
        return
    gb = B * world
    print(json.dumps({
        "impl": "reference", "metric": "images_per_sec", "value": gb / ms * 1e3, "unit": "img/s", "n_gpus": world,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "fp32 (cuDNN TF32 conv, fp32 cuBLAS fc)", "data": "synthetic",
        "config": {"model": "reference ConvNet (baseline/_ref/mnist_distributed.py), SGD lr 1e-4, torch DDP + NCCL",
                   "global_batch": gb, "seq_len": None, "image": [S, S], "per_gpu_batch": B,
                   "parallelism": f"dp{world}", "l2": "working set per step >> 126 MB L2; no flush needed"},
        "clocks": {"sm_mhz": clocks["sm_mhz"], "sm_max_mhz": clocks["sm_max_mhz"], "reasons": clocks["reasons"]},
        "e2e": {"value": gb / ms_e2e * 1e3, "unit": "img/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": int(B * S * S * 4 + B * 8), "d2h_bytes_per_step": 4,
                "note": "pre-resized fp32 batch in pinned host memory (the reference's CPU resize is not charged)"},
        "e2e_stock_loader": stock, "gpu_launches": None, "final_loss": final_loss,
        "peak_mem_gib": round(peak, 2),
    }))


if __name__ == "__main__":
    if len(sys.argv) >= 6 and sys.argv[1] == "--stock":
        _stock_main(sys.argv[2:])
