"""One small training step (every kernel of the engine once, tcgen05 paths included) - the target of
compute-sanitizer runs.  usage: sanitize_step.py [H W B]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from torch_distributed_sandbox_b200.models.engine import FusedConvNet

H, W, B = (int(a) for a in (sys.argv[1:4] if len(sys.argv) >= 4 else (200, 264, 2)))
eng = FusedConvNet((H, W), batch_size=B, lr=1e-3, fuse_fc_sgd=True)
g = torch.Generator().manual_seed(0)
imgs = torch.randint(0, 256, (B, 28, 28), dtype=torch.uint8, generator=g).cuda()
labels = torch.randint(0, 10, (B,), generator=g).cuda()
for _ in range(2):
    loss = eng.train_step(imgs, labels)
torch.cuda.synchronize()
print(f"sanitize_step {H}x{W} batch {B}: loss {float(loss):.5f} ok")
