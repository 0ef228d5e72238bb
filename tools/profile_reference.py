"""Where do the reference's 800 ms per step go?  torch.profiler over two steps of the
unmodified reference ConvNet + DDP + SGD (baseline/_ref), top CUDA kernels by device time.
Output: gpurun_out/reference_profile.txt (summarised in BASELINE.md)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
import torch
import torch.distributed as dist
import torch.nn as nn
from torch.profiler import ProfilerActivity, profile

import mnist_distributed as ref

os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", ref.find_free_port())
dist.init_process_group("nccl", rank=0, world_size=1)
torch.manual_seed(0)
model = ref.ConvNet()
model(torch.zeros([1, 1] + ref.IMAGE_SHAPE))
torch.cuda.set_device(0)
model.cuda(0)
crit = nn.CrossEntropyLoss().cuda(0)
opt = torch.optim.SGD(model.parameters(), 1e-4)
model = nn.parallel.DistributedDataParallel(model, device_ids=[0])
x = torch.rand(5, 1, 3000, 3000, device="cuda")
y = torch.randint(0, 10, (5,), device="cuda")


def step():
    out = model(x)
    loss = crit(out, y)
    opt.zero_grad()
    loss.backward()
    opt.step()


for _ in range(3):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(2):
        step()
    torch.cuda.synchronize()
os.makedirs("gpurun_out", exist_ok=True)
txt = prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=90)
open("gpurun_out/reference_profile.txt", "w").write(txt)
print(txt[-6000:])
dist.destroy_process_group()
