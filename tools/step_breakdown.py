"""Per-kernel CUDA-event times of one full-size training step on ONE GPU.
usage: step_breakdown.py [--precision bf16|tf32] [--ver N] [--batch B] [--image S]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from torch_distributed_sandbox_b200 import ops
from torch_distributed_sandbox_b200.models.engine import FusedConvNet

ap = argparse.ArgumentParser()
ap.add_argument("--precision", default="bf16")
ap.add_argument("--ver", type=int, default=5)
ap.add_argument("--batch", type=int, default=5)
ap.add_argument("--image", type=int, default=3000)
a = ap.parse_args()
ops.conv_tc_set_version(a.ver)
B = a.batch
eng = FusedConvNet((a.image, a.image), batch_size=B, lr=1e-4, precision=a.precision)
g = torch.Generator().manual_seed(0)
imgs = torch.randint(0, 256, (B, 28, 28), dtype=torch.uint8, generator=g).cuda()
labels = torch.randint(0, 10, (B,), generator=g).cuda()
for _ in range(3):
    eng.train_step(imgs, labels)
torch.cuda.synchronize()
skip = ("sm_count", "l2_flush", "set_reserved_sms", "c_void_p", "c_int", "c_float", "c_double", "c_size_t")
names = [n for n in dir(ops) if callable(getattr(ops, n)) and not n.startswith("_") and n not in skip
         and not n.endswith("_grid") and not n.startswith("conv_tc_")]
times, orig = {}, {n: getattr(ops, n) for n in names}


def wrap(n):
    f = orig[n]

    def g_(*args, **kw):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        r = f(*args, **kw)
        e.record()
        times.setdefault(n, []).append((s, e))
        return r
    return g_


for n in names:
    setattr(ops, n, wrap(n))
acc = {}
for it in range(6):
    times.clear()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    eng.train_step(imgs, labels)
    t1.record()
    torch.cuda.synchronize()
    if it >= 2:
        for n, evs in times.items():
            acc.setdefault(n, []).append(sum(s.elapsed_time(e) for s, e in evs))
        acc.setdefault("_step(with event overhead)", []).append(t0.elapsed_time(t1))
for n in names:
    setattr(ops, n, orig[n])
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(20):
    eng.train_step(imgs, labels)
e.record()
torch.cuda.synchronize()
print(f"--- 1 GPU, precision {a.precision}, tc v{a.ver}, batch {B}, {a.image}^2: {s.elapsed_time(e) / 20:.3f} ms/step "
      f"(20 steps back to back); loss {float(eng.loss):.4f}; per-kernel mean ms over 4 steps:")
for n, v in sorted(acc.items(), key=lambda kv: -sum(kv[1])):
    print(f"  {n:28s} {sum(v) / len(v):8.3f}")
