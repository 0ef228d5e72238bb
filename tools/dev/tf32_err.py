import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.nn.functional as F
from torch_distributed_sandbox_b200.models import reference_init
from torch_distributed_sandbox_b200.models.engine import FusedConvNet
from torch_distributed_sandbox_b200.data.mnist import synthetic_digits
DEV = "cuda:0"
def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-12))
for shape, B in (((64, 96), 4), ((256, 384), 4), ((1000, 1000), 2)):
    imgs, labs = synthetic_digits(B, seed=5)
    imgs = torch.from_numpy(imgs).view(B, 28, 28).to(DEV); labs = torch.from_numpy(labs.astype("int64")).to(DEV)
    res = {}
    for mode in ("fp32", "cudnn_tf32", "ours_tf32", "ours_bf16"):
        if mode.startswith("ours"):
            eng = FusedConvNet(shape, batch_size=B, device=DEV, lr=1e-3, fuse_fc_sgd=False, precision=mode[5:])
            eng.forward(imgs); eng.loss_backward(labs)
            res[mode] = {k: v.clone() for k, v in eng.named_gradients().items()}
            res[mode]["loss"] = eng.loss.clone()
        else:
            torch.backends.cudnn.allow_tf32 = mode == "cudnn_tf32"
            torch.backends.cuda.matmul.allow_tf32 = False
            ref = reference_init(shape, seed=0).to(DEV)
            e0 = FusedConvNet(shape, batch_size=B, device=DEV, precision="tf32"); e0.set_input(imgs)
            x = e0.x[:, None].clone()
            loss = F.cross_entropy(ref(x), labs); loss.backward()
            res[mode] = {k: p.grad.clone() for k, p in ref.named_parameters()}
            res[mode]["loss"] = loss.detach()
    print(f"--- shape {shape} B {B}: rel-L2 vs torch fp32 (TF32 off)")
    for k in res["fp32"]:
        if k in ("layer1.0.bias", "layer2.0.bias"): continue
        print(f"  {k:18s} cudnn_tf32 {rel(res['cudnn_tf32'][k], res['fp32'][k]):.3e}   ours_tf32 {rel(res['ours_tf32'][k], res['fp32'][k]):.3e}   ours_bf16 {rel(res['ours_bf16'][k], res['fp32'][k]):.3e}")
