"""Loss curves at the flagship size: the UNMODIFIED reference ConvNet (baseline/_ref/mnist_onegpu.py, fp32, TF32 off:
the exact oracle) vs the engine in "tf32" mode (the reference's permitted numerics) vs the engine in "bf16" mode.
Same seed-0 initial state, same non-repeating synthetic digits, same SGD(1e-4), batch 5, 3000x3000.
usage: convergence.py [steps] [image] [lr] -> gpurun_out/convergence_<image>_lr<lr>.txt
The reference's lr = 1e-4 puts this model (18M-input linear layer) in a violently unstable regime: the loss jumps from 2.2
to 50-90 within two steps, so trajectories agree for a few steps and then separate chaotically (all reach ~0 by step 40).
A smaller lr gives a stable regime in which the curves can be compared step by step."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
import torch
import torch.nn as nn

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
from torch_distributed_sandbox_b200.data.mnist import synthetic_digits
from torch_distributed_sandbox_b200.models.engine import FusedConvNet

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
S = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
LR = float(sys.argv[3]) if len(sys.argv) > 3 else 1e-4
B, DEV = 5, "cuda:0"
import mnist_onegpu as ref

ref.IMAGE_SHAPE = [S, S]
torch.manual_seed(0)
model = ref.ConvNet()
model(torch.zeros([1, 1] + ref.IMAGE_SHAPE))            # the reference's lazy init (mnist_onegpu.py:39)
sd0 = {k: v.clone() for k, v in model.state_dict().items()}
model.cuda(0)
crit = nn.CrossEntropyLoss().cuda(0)
opt = torch.optim.SGD(model.parameters(), LR)
engs = {p: FusedConvNet((S, S), batch_size=B, device=DEV, lr=LR, init="empty", precision=p) for p in ("tf32", "bf16")}
for e in engs.values():
    e.load_state_dict(sd0)
imgs, labs = synthetic_digits(steps * B, seed=4242)
imgs = torch.from_numpy(imgs).view(steps, B, 28, 28).to(DEV)
labs = torch.from_numpy(labs.astype("int64")).view(steps, B).to(DEV)
curves = {"ref_fp32": [], "tf32": [], "bf16": []}
for s in range(steps):
    engs["tf32"].set_input(imgs[s])
    x = engs["tf32"].x[:, None].clone()                  # the PIL-exact resized batch, identical for all three
    loss = crit(model(x), labs[s])
    opt.zero_grad()
    loss.backward()
    opt.step()
    curves["ref_fp32"].append(float(loss.detach()))
    for p, e in engs.items():
        curves[p].append(float(e.train_step(imgs[s], labs[s])))
torch.cuda.synchronize()


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-12))


out = []
out.append(f"# {steps} SGD({LR:g}) steps, batch {B}, {S}x{S}, non-repeating synthetic digits (seed 4242), seed-0 reference init")
out.append("# step   reference_fp32   engine_tf32   engine_bf16")
for s in range(steps):
    if s < 20 or s % 10 == 0 or s == steps - 1:
        out.append(f"{s:5d}   {curves['ref_fp32'][s]:12.5f}   {curves['tf32'][s]:12.5f}   {curves['bf16'][s]:12.5f}")
import statistics
for lo, hi in ((0, 10), (10, 50), (50, 100), (100, steps)):
    if lo >= steps:
        break
    hi = min(hi, steps)
    r = statistics.mean(curves["ref_fp32"][lo:hi])
    out.append(f"# mean loss steps [{lo},{hi}): reference {r:.5f}  tf32 {statistics.mean(curves['tf32'][lo:hi]):.5f}  "
               f"bf16 {statistics.mean(curves['bf16'][lo:hi]):.5f}")
    for p in ("tf32", "bf16"):
        d = max(abs(a - b) for a, b in zip(curves[p][lo:hi], curves["ref_fp32"][lo:hi]))
        out.append(f"#   max |loss_{p} - loss_ref| in [{lo},{hi}): {d:.5f}")
sd_ref = model.state_dict()
for p, e in engs.items():
    sd = e.state_dict()
    out.append(f"# final state_dict distance (rel-L2) {p} vs reference: " +
               "  ".join(f"{k}={rel(sd[k], sd_ref[k]):.2e}" for k in sd_ref if not k.endswith("num_batches_tracked")))
text = "\n".join(out)
print(text)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
open(os.path.join(ROOT, "gpurun_out", f"convergence_{S}_lr{LR:g}.txt"), "w").write(text + "\n")
