"""Run ONE kernel of the training step at the flagship size (batch 5, 3000x3000) a few times,
standalone, so it can be profiled with `ncu -k regex:<name>` (B200_PROFILING.md recipe).
usage: run_one_kernel.py <conv2_fwd|conv2_dgrad|conv2_dgrad_stats|conv2_wgrad|conv1_fwd|conv1_wgrad|step> [tc version] [iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from torch_distributed_sandbox_b200 import ops
from torch_distributed_sandbox_b200.models.engine import FusedConvNet

which = sys.argv[1]
ver = int(sys.argv[2]) if len(sys.argv) > 2 else 5
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
ops.conv_tc_set_version(ver)
B = 5
eng = FusedConvNet((3000, 3000), batch_size=B, lr=1e-4, fuse_fc_sgd=True)
g = torch.Generator().manual_seed(0)
imgs = torch.randint(0, 256, (B, 28, 28), dtype=torch.uint8, generator=g).cuda()
labels = torch.randint(0, 10, (B,), generator=g).cuda()
eng.train_step(imgs, labels)          # populates every activation / gradient buffer with real data
eng.forward(imgs)
torch.cuda.synchronize()
P = eng.params
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
fn = {
    "conv2_fwd": lambda: ops.conv2_fwd_tc(eng.p1, P["layer2.0.weight"], P["layer2.0.bias"], eng.x2, eng.part_c2),
    "conv2_dgrad": lambda: ops.conv2_dgrad_tc(eng.x2, P["layer2.0.weight"], torch.empty_like(eng.p1)),
    "conv2_dgrad_stats": lambda: ops.conv2_dgrad_stats_tc(eng.dx2, P["layer2.0.weight"], eng.p1, eng.bn1,
                                                          torch.empty_like(eng.p1), eng.part_bwd),
    "conv2_wgrad": lambda: ops.conv2_wgrad_tc(eng.p1, eng.x2, eng.part_wg2, eng.grads["layer2.0.weight"]),
    "conv1_fwd_tc": lambda: ops.conv1_fwd_tc(eng.x, P["layer1.0.weight"], P["layer1.0.bias"], eng.x1, eng.part_c1),
    "conv1_fwd": lambda: ops.conv1_fwd(eng.x, P["layer1.0.weight"], P["layer1.0.bias"], eng.x1, eng.part_c1),
    "conv1_wgrad_tc": lambda: ops.conv1_wgrad_tc(eng.x, eng.x1, eng.p1, eng.bn1, eng.red1, eng.part_wg1,
                                                 eng.grads["layer1.0.weight"]),
    "conv1_wgrad": lambda: ops.conv1_wgrad(eng.x, eng.x1, eng.p1, eng.bn1, eng.red1, eng.part_wg1,
                                           eng.grads["layer1.0.weight"]),
    "step": lambda: eng.train_step(imgs, labels),
}[which]
fn()
torch.cuda.synchronize()
s.record()
for _ in range(iters):
    fn()
e.record()
torch.cuda.synchronize()
print(f"{which} (tc v{ver}): {s.elapsed_time(e) / iters:.3f} ms")
