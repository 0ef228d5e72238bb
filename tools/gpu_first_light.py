"""First-light script for a B200 box: per-kernel device timings of one full-size training
step (3000x3000, batch 5), written to gpurun_out/first_light.txt."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from torch_distributed_sandbox_b200 import ops
from torch_distributed_sandbox_b200.models.engine import FusedConvNet


def main():
    os.makedirs("gpurun_out", exist_ok=True)
    out = open("gpurun_out/first_light.txt", "w")

    def log(*a):
        s = " ".join(str(x) for x in a)
        print(s)
        out.write(s + "\n")
        out.flush()

    backend = sys.argv[1] if len(sys.argv) > 1 else "tc"
    if backend in ("tc1", "tc2", "tc3"):
        ops.conv_tc_set_version(int(backend[2]))
        backend = "tc"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    t0 = time.time()
    eng = FusedConvNet((3000, 3000), batch_size=B, lr=1e-4, conv_backend=backend)
    torch.cuda.synchronize()
    log(f"engine built in {time.time() - t0:.1f}s; activations {eng.activation_bytes() / 2**30:.2f} GiB; "
        f"max mem {torch.cuda.max_memory_allocated() / 2**30:.2f} GiB")
    g = torch.Generator().manual_seed(0)
    imgs = torch.randint(0, 256, (B, 28, 28), dtype=torch.uint8, generator=g).cuda()
    labels = torch.randint(0, 10, (B,), generator=g).cuda()
    for i in range(3):
        loss = eng.train_step(imgs, labels)
    torch.cuda.synchronize()
    log("warm loss", float(loss))

    # patch every op with event timing
    names = [n for n in dir(ops) if callable(getattr(ops, n)) and not n.startswith("_")
             and n not in ("sm_count", "l2_flush") and not n.endswith("_grid")]
    times = {}
    orig = {n: getattr(ops, n) for n in names}

    def wrap(n):
        f = orig[n]

        def g_(*a, **k):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            r = f(*a, **k)
            e.record()
            times.setdefault(n, []).append((s, e))
            return r
        return g_

    for n in names:
        setattr(ops, n, wrap(n))
    for i in range(3):
        times.clear()
        eng.train_step(imgs, labels)
        torch.cuda.synchronize()
    tot = 0.0
    for n, evs in times.items():
        ms = [s.elapsed_time(e) for s, e in evs]
        tot += sum(ms)
        log(f"  {n:24s} x{len(ms)}  {sum(ms):8.3f} ms   ({', '.join(f'{m:.3f}' for m in ms)})")
    log(f"  sum of kernels: {tot:.3f} ms")
    for n in names:
        setattr(ops, n, orig[n])
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    s.record()
    K = 10
    for i in range(K):
        eng.train_step(imgs, labels)
    e.record()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e) / K
    log(f"train_step: {ms:.3f} ms/step -> {B / ms * 1e3:.1f} img/s  (loss {float(eng.loss):.4f})")
    log(f"max mem {torch.cuda.max_memory_allocated() / 2**30:.2f} GiB")


if __name__ == "__main__":
    main()
