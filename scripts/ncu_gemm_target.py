"""ncu target: the tcgen05 3xTF32 GEMM on the layer-2 forward shape (plain and implicit-conv producer)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch_distlearn_b200.ops import tc_gemm
from torch_distlearn_b200.ops.conv import _tc_conv
M, N, K = 32768, 128, 1600
a = torch.randn(M, K, device="cuda"); b = torch.randn(N, K, device="cuda"); c = torch.empty(M, N, device="cuda")
x = torch.randn(128, 16, 16, 64, device="cuda")
for _ in range(3):
    tc_gemm(a, b, out=c, terms=3)
    _tc_conv(1, x, b, c, None, 128, 16, 16, 64, 128, 1, 3)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); tc_gemm(a, b, out=c, terms=3); e1.record(); torch.cuda.synchronize()
print("plain gemm us", e0.elapsed_time(e1) * 1e3)
