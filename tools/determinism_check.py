"""Run the same training steps on two identically initialised engines and compare every buffer
bit for bit (race / nondeterminism detector for the hand-written kernels)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch_distributed_sandbox_b200.models.engine import FusedConvNet

shape = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (96, 200)
B = 3
g = torch.Generator().manual_seed(0)
x = torch.rand(4, B, 1, *shape, generator=g).cuda()
y = torch.randint(0, 10, (4, B), generator=g).cuda()
bad = 0
for trial in range(3):
    e1 = FusedConvNet(shape, batch_size=B, lr=1e-3, fuse_fc_sgd=False)
    e2 = FusedConvNet(shape, batch_size=B, lr=1e-3, fuse_fc_sgd=False)
    for s in range(4):
        for e in (e1, e2):
            e.forward(x[s]); e.loss_backward(y[s])
        torch.cuda.synchronize()
        for name in ("x1", "p1", "x2", "p2", "small_grad", "fcw_grad", "bn1", "bn2", "red1", "red2", "loss"):
            a, b = getattr(e1, name), getattr(e2, name)
            if not torch.equal(a, b):
                d = (a.float() - b.float()).abs().max().item()
                print(f"trial {trial} step {s}: {name} differs (max abs {d:.3e})")
                bad += 1
        for e in (e1, e2):
            e.step()
print("deterministic" if bad == 0 else f"{bad} mismatches")
