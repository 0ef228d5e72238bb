"""The reference's one experiment (README.md:9-15): at 3000x3000, batch 10 does not fit a 24 GB GPU,
batch 5 does, and data parallelism restores the effective batch.  Reproduced here on a B200 by capping
the process to 24 GB (torch.cuda.set_per_process_memory_fraction), for the unmodified reference model and
for the fused engine; then the largest per-GPU batch each can hold in the full 180 GB.
Output: gpurun_out/oom_repro.txt"""
import gc
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
import torch

out = open(os.path.join(ROOT, "gpurun_out", "oom_repro.txt"), "w") if os.path.isdir(os.path.join(ROOT, "gpurun_out")) else sys.stdout


def log(*a):
    s = " ".join(str(x) for x in a)
    print(s, flush=True)
    if out is not sys.stdout:
        out.write(s + "\n")
        out.flush()


def _fresh():
    """Start every measurement from an empty pool and a reset peak counter (an OOM's traceback keeps its tensors
    alive until the handler has returned, so the clean-up in `finally` alone left a stale peak for the next run)."""
    gc.collect()
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()


def try_reference(batch, steps=2):
    import torch.nn as nn
    import mnist_onegpu as ref
    _fresh()
    try:
        torch.manual_seed(0)
        model = ref.ConvNet()
        model(torch.zeros([1, 1] + ref.IMAGE_SHAPE))
        model.cuda(0)
        crit = nn.CrossEntropyLoss().cuda(0)
        opt = torch.optim.SGD(model.parameters(), 1e-4)
        x = torch.rand(batch, 1, 3000, 3000, device="cuda")
        y = torch.randint(0, 10, (batch,), device="cuda")
        for _ in range(steps):
            loss = crit(model(x), y)
            opt.zero_grad()
            loss.backward()
            opt.step()
        torch.cuda.synchronize()
        return True, torch.cuda.max_memory_allocated() / 2 ** 30
    except torch.OutOfMemoryError:
        return False, torch.cuda.max_memory_allocated() / 2 ** 30
    finally:
        gc.collect()
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()


def try_engine(batch, steps=3, timeit=False, precision="bf16"):
    from torch_distributed_sandbox_b200.models.engine import FusedConvNet
    eng = None
    _fresh()
    try:
        eng = FusedConvNet((3000, 3000), batch_size=batch, lr=1e-4, seed=0, precision=precision)
        g = torch.Generator().manual_seed(0)
        imgs = torch.randint(0, 256, (batch, 28, 28), dtype=torch.uint8, generator=g).cuda()
        labels = torch.randint(0, 10, (batch,), generator=g).cuda()
        for _ in range(steps):
            eng.train_step(imgs, labels)
        torch.cuda.synchronize()
        ms = None
        if timeit:
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(3):
                eng.train_step(imgs, labels)
            e.record()
            torch.cuda.synchronize()
            ms = s.elapsed_time(e) / 3
        return True, torch.cuda.max_memory_allocated() / 2 ** 30, ms
    except torch.OutOfMemoryError:
        return False, torch.cuda.max_memory_allocated() / 2 ** 30, None
    finally:
        del eng
        gc.collect()
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()


def main():
    total = torch.cuda.get_device_properties(0).total_memory
    log(f"device: {torch.cuda.get_device_name(0)}, {total / 2 ** 30:.1f} GiB")
    only = os.environ.get("OOM_ONLY")          # "tf32": just the engine's tf32 rows (a cheap refresh of those lines)
    torch.cuda.set_per_process_memory_fraction(min(1.0, 24e9 / total), 0)
    log("== capped at 24 GB (the RTX A5000 of README.md:11)")
    if not only:
        for b in (5, 10):
            ok, peak = try_reference(b)
            log(f"reference  batch {b:3d}: {'fits' if ok else 'OOM '}  peak {peak:5.1f} GiB")
    for prec in (("bf16", "tf32") if not only else (only,)):
        for b in (5, 10, 20, 36):
            ok, peak, _ = try_engine(b, precision=prec)
            log(f"b200dp {prec} batch {b:3d}: {'fits' if ok else 'OOM '}  peak {peak:5.1f} GiB")
    torch.cuda.set_per_process_memory_fraction(1.0, 0)
    log("== full device")
    if not only:
        for b in (32, 48, 64):
            ok, peak = try_reference(b, steps=1)
            log(f"reference  batch {b:3d}: {'fits' if ok else 'OOM '}  peak {peak:5.1f} GiB")
            if not ok:
                break
    for prec, batches in ((("bf16", (64, 128, 256, 288)), ("tf32", (64, 128, 256, 288))) if not only
                          else ((only, (64, 128, 256, 288)),)):
        for b in batches:
            ok, peak, ms = try_engine(b, timeit=True, precision=prec)
            log(f"b200dp {prec} batch {b:3d}: {'fits' if ok else 'OOM '}  peak {peak:5.1f} GiB" +
                (f"  {ms:.1f} ms/step = {b / ms * 1e3:.0f} img/s" if ms else ""))
            if not ok:
                break


if __name__ == "__main__":
    main()
