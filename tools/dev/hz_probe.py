"""Time (CUDA events) the two layer-1 forward passes of the tf32 mode at full size: row-operand kernels vs the bf16 hi/lo ones."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from torch_distributed_sandbox_b200 import ops

B, S = 5, 3000
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 20
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(1)
x = torch.rand(B, S, S, device=dev, generator=g)
w = torch.randn(16, 1, 5, 5, device=dev, generator=g) * 0.2
b = torch.randn(16, device=dev, generator=g) * 0.1
bn = torch.zeros(4, 16, device=dev); bn[0] = 1.0
part = torch.zeros(ops.conv1_fwd_grid(), 32, device=dev)
p1 = torch.empty(B, S // 2, S // 2, 16, device=dev)
mask = torch.empty(B, S // 2, S // 2, 16, device=dev, dtype=torch.uint8)


def timeit(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


for ro in (True, False):
    t1 = timeit(lambda: ops.conv1_stats_tc(x, w, b, part, row_operand=ro))
    t2 = timeit(lambda: ops.conv1_fwd_pool_tc(x, w, b, bn, None, p1, part, amask_u8=mask, row_operand=ro))
    print(f"row_operand={ro}: stats {t1:.3f} ms, pool+mask {t2:.3f} ms")
big = torch.empty(B, S // 2, S // 2, 32, device=dev)
t = timeit(lambda: p1.fill_(1.0)); print(f"fill 720 MB: {t:.3f} ms = {0.72 / t:.2f} TB/s")
t = timeit(lambda: big.fill_(1.0)); print(f"fill 1440 MB: {t:.3f} ms = {1.44 / t:.2f} TB/s")
t = timeit(lambda: p1.copy_(big[..., :16].contiguous() if False else p1)); 
src = torch.empty_like(p1)
t = timeit(lambda: p1.copy_(src)); print(f"copy 720 MB -> 720 MB: {t:.3f} ms = {1.44 / t:.2f} TB/s")
