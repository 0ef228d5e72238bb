import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from torch_distributed_sandbox_b200 import ops
torch.backends.cudnn.allow_tf32 = False
DEV = "cuda:0"
torch.set_printoptions(linewidth=200, precision=3, sci_mode=False)
for dt in (torch.float32, torch.bfloat16):
    B, H, W = 1, 8, 64
    x = torch.zeros(B, H, W, 16, device=DEV, dtype=dt)
    dy = torch.zeros(B, H, W, 32, device=DEV, dtype=dt)
    x[..., :] = 1.0
    dy[..., :] = 1.0
    part = torch.zeros(ops.conv_tc_max_grid(), 32 * 16 * 25, device=DEV)
    dw = torch.full((32, 16, 5, 5), -7.0, device=DEV)
    ops.conv2_wgrad_tc(x, dy, part, dw)
    torch.cuda.synchronize()
    print(dt, "dw[0,0]=\n", dw[0, 0], "\ndw[31,15]=\n", dw[31, 15], "\nabs sum", float(dw.abs().sum()), "nonzero", int((dw != 0).sum()))
    ref = torch.nn.grad.conv2d_weight(x.float().permute(0, 3, 1, 2), (32, 16, 5, 5), dy.float().permute(0, 3, 1, 2), padding=2)
    print("ref[0,0]=\n", ref[0, 0])
    print("partials row0 nonzero:", int((part[0] != 0).sum()), "max", float(part[0].abs().max()))
