"""One steady-state full-size training step between cudaProfilerStart/Stop, for
`ncu --profile-from-start off --set full`: every kernel of the step is captured once.
usage: ncu_step.py [bf16|tf32] [batch]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from torch_distributed_sandbox_b200.models.engine import FusedConvNet

prec = sys.argv[1] if len(sys.argv) > 1 else "tf32"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 5
eng = FusedConvNet((3000, 3000), batch_size=B, lr=1e-4, precision=prec)
g = torch.Generator().manual_seed(0)
imgs = torch.randint(0, 256, (B, 28, 28), dtype=torch.uint8, generator=g).cuda()
labels = torch.randint(0, 10, (B,), generator=g).cuda()
for _ in range(3):
    eng.train_step(imgs, labels)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
eng.train_step(imgs, labels)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print(f"step ok ({prec}, batch {B}): loss {float(eng.loss):.4f}")
