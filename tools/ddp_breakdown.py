"""Per-kernel device times of the DDP training step (run under torch.distributed.run): shows which
main-stream kernels stretch while the bucket all-reduce runs on the side stream."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from torch_distributed_sandbox_b200 import distributed as dist, ops
from torch_distributed_sandbox_b200.models.module import FusedModule
from torch_distributed_sandbox_b200.parallel import DistributedDataParallel

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
B = 5
model = DistributedDataParallel(FusedModule((3000, 3000), batch_size=B, device=dev, lr=1e-4), device_ids=[local])
g = torch.Generator().manual_seed(rank)
imgs = torch.randint(0, 256, (B, 28, 28), dtype=torch.uint8, generator=g).to(dev)
labels = torch.randint(0, 10, (B,), generator=g).to(dev)
for _ in range(5):
    model.train_step(imgs, labels)
torch.cuda.synchronize()
names = [n for n in dir(ops) if callable(getattr(ops, n)) and not n.startswith("_") and n not in ("sm_count", "l2_flush")
         and not n.endswith("_grid") and not n.startswith("conv_tc_") and n not in ("c_void_p", "c_int", "c_float", "c_double", "c_size_t")]
times, orig = {}, {n: getattr(ops, n) for n in names}


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
acc = {}
for it in range(6):
    times.clear()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    model.train_step(imgs, labels)
    t1.record()
    torch.cuda.synchronize()
    if it >= 2:
        for n, evs in times.items():
            acc.setdefault(n, []).append(sum(s.elapsed_time(e) for s, e in evs))
        acc.setdefault("_step", []).append(t0.elapsed_time(t1))
if rank == 0:
    tag = "skipfc" if os.environ.get("B200DP_DEBUG_SKIP_FC_ALLREDUCE") == "1" else "allreduce"
    print(f"--- DDP world {world} ({tag}) mean ms over 4 steps")
    for n, v in sorted(acc.items(), key=lambda kv: -sum(kv[1])):
        print(f"  {n:24s} {sum(v) / len(v):8.3f}")
dist.destroy_process_group()
