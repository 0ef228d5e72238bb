import os, sys, ctypes
from ctypes import c_void_p, c_int
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from torch_distributed_sandbox_b200 import ops, _native
lib = _native.cuda(); ops._lib()
B, S = 5, 3000
x = torch.rand(B, S, S, device="cuda"); w = torch.randn(16, 1, 5, 5, device="cuda") * 0.2; b = torch.zeros(16, device="cuda")
x1 = torch.empty(B, S, S, 16, device="cuda"); part = torch.zeros(ops.conv1_fwd_grid(), 32, device="cuda")
n = c_int(0)
def run(ptr):
    return lib.b200_conv1_fwd_tc_f32(c_void_p(x.data_ptr()), c_void_p(w.data_ptr()), c_void_p(b.data_ptr()), c_void_p(ptr), c_void_p(part.data_ptr()), B, S, S, ctypes.byref(n), c_void_p(torch.cuda.current_stream().cuda_stream))
for name, ptr in (("store", x1.data_ptr()), ("stats-only", 0)):
    for _ in range(3): run(ptr)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(10): run(ptr)
    e.record(); torch.cuda.synchronize()
    print(name, s.elapsed_time(e) / 10, "ms", float(part[:n.value].sum()))
