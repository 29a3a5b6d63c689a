"""ncu target: the three implicit GEMMs of layer 3 (forward, dgrad, wgrad), one launch each after warm-up — to see why the
weight-gradient GEMM runs at 125 TFLOP/s where forward and dgrad reach 200 (profiles/tail_probe.md).

    python scripts/ncu_gemm_modes_target.py && ncu --set full --clock-control none --import-source on \
        -k regex:tc_gemm -s 9 -c 3 -o gpurun_out/prof_gemm_modes -f python scripts/ncu_gemm_modes_target.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from torch_distlearn_b200.ops import conv as convops

B, H, cin, cout = 128, 8, 128, 256
M = B * H * H
dev = "cuda"
x = torch.randn(B, H, H, cin, device=dev)
w = torch.randn(cout, 25 * cin, device=dev) * 0.05
bias = torch.randn(cout, device=dev)
dy = torch.randn(B, H, H, cout, device=dev)
yo = torch.empty(M, cout, device=dev)
dxo = torch.empty(B, H, H, cin, device=dev)
dwo = torch.zeros(cout, 25 * cin, device=dev)
sp_f = convops._splits((M + 127) // 128, (cout + 127) // 128, 25 * cin)
sp_d = convops._splits((M + 127) // 128, (cin + 127) // 128, 25 * cout)
sp_w = convops._splits((cout + 127) // 128, (25 * cin + 127) // 128, M)


def run():
    convops._tc_conv(1, x, w, yo, bias, B, H, H, cin, cout, sp_f, 3)
    convops._tc_conv(2, dy, w, dxo, None, B, H, H, cin, cout, sp_d, 3)
    convops._tc_conv(3, x, dy, dwo, None, B, H, H, cin, cout, sp_w, 3, accumulate=True)


for _ in range(3):      # 9 launches of warm-up: ncu skips them (-s 9)
    run()
torch.cuda.synchronize()
e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
e[0].record()
convops._tc_conv(1, x, w, yo, bias, B, H, H, cin, cout, sp_f, 3)
e[1].record()
convops._tc_conv(2, dy, w, dxo, None, B, H, H, cin, cout, sp_d, 3)
e[2].record()
convops._tc_conv(3, x, dy, dwo, None, B, H, H, cin, cout, sp_w, 3, accumulate=True)
e[3].record()
torch.cuda.synchronize()
print("forward %.1f us, dgrad %.1f us, wgrad %.1f us" % tuple(e[i].elapsed_time(e[i + 1]) * 1e3 for i in range(3)))
