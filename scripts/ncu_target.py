#!/usr/bin/env python
"""Small single-GPU program for `ncu`: launches the fused all-reduce+SGD kernel (a) as world=1 on a
256 MiB vector (local HBM-bound path) and (b) as 4 emulated ranks in one cooperative launch on
4 x 64 MiB vectors (the real multi-rank code path with peers in local memory), plus the fused
elastic-averaging kernel the same way.  Prints CUDA-event timings so the plain run is informative too."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import torch_distlearn_b200 as dl

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)


def timed(fn, iters=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


n1 = 64 << 20                         # 256 MiB of fp32
c1 = dl.P2PCommunicator(rank=0, world=1, device=dev, heap_mb=3 * 256 + 64)
g, p, m = c1.alloc(n1), c1.alloc(n1), c1.alloc(n1)
g.tensor.normal_(); p.tensor.normal_()
us = timed(lambda: c1.fused_sgd(g, p, m, n1, 1e-3, momentum=0.9, active=1, zero_grads=True))
# bytes: read g, p, m; write p, m, g(zero)  = 6 * 4 * n
print(f"world=1 fused_sgd+momentum+zero 256MiB: {us:.1f} us  -> {6 * 4 * n1 / us / 1e3:.0f} GB/s of HBM traffic")
us = timed(lambda: c1.fused_sgd(g, p, None, n1, 1e-3, active=1, zero_grads=True))
print(f"world=1 fused_sgd+zero 256MiB: {us:.1f} us  -> {4 * 4 * n1 / us / 1e3:.0f} GB/s of HBM traffic")

W, n4 = 4, 16 << 20                   # 4 emulated ranks x 64 MiB
ce = dl.EmulatedCommunicator(W, dev, heap_mb=4 * 64 + 64)
ge, gv = ce.alloc_all(n4)
pe, pv = ce.alloc_all(n4)
x0, _ = ce.alloc_all(n4)
x1, _ = ce.alloc_all(n4)
for v in gv:
    v.normal_()
us = timed(lambda: ce.fused_sgd(ge, pe, None, n4, 1e-3, active=(1 << W) - 1, zero_grads=True, variant="twoshot"))
print(f"emulated world=4 fused_sgd twoshot 4x64MiB: {us:.1f} us")
cur = [0]


def ea():
    a, b = (x0, x1) if cur[0] == 0 else (x1, x0)
    ce.fused_ea(pe, a, b, n4, 0.1, active=(1 << W) - 1, variant="twoshot")
    cur[0] ^= 1


us = timed(ea)
print(f"emulated world=4 fused_ea twoshot 4x64MiB: {us:.1f} us")
