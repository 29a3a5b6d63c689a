#!/usr/bin/env python
"""Single-GPU program for `ncu`: one or two launches of every kernel family that had no capture in round 1 — the
second-generation collectives as 8 emulated ranks in ONE cooperative launch (the real multi-rank code path with the peers'
heaps in local memory; ncu may not wrap a multi-process job, B200_PROFILING.md), and the tail kernels of the flagship
model at its layer-2 shape.  Prints CUDA-event timings so that the plain run is informative too.

    python scripts/ncu_kernels_target.py && \
    ncu --set full --clock-control none --import-source on \
        -k regex:"pipe_kernel|ll_kernel|broadcast_wide|allreduce_twoshot|async_ea_client|bn_|colsum|head_|im2col" \
        -c 40 -o gpurun_out/prof_kernels python scripts/ncu_kernels_target.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import torch_distlearn_b200 as dl
from torch_distlearn_b200 import _lib

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
lib = _lib.cuda()


def timed(name, fn, iters=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:58s} {e0.elapsed_time(e1) / iters * 1e3:9.1f} us", flush=True)


# ---- collectives, 8 emulated ranks, 8 x 16 MiB (the flagship model's gradient size) and a small message ----------
W, n = 8, 4 << 20
ce = dl.EmulatedCommunicator(W, dev, heap_mb=5 * 16 + 48)
full = (1 << W) - 1
g, gv = ce.alloc_all(n)
p, _ = ce.alloc_all(n)
c0, _ = ce.alloc_all(n)
c1, _ = ce.alloc_all(n)
for v in gv:
    v.normal_()
timed("fused_sgd hybrid (all P2P) 8 x 16 MiB", lambda: ce.fused_sgd(g, p, None, n, 1e-3, active=full, zero_grads=True, variant="hybrid"))
timed("allreduce hybrid (all P2P) 8 x 16 MiB", lambda: ce.allreduce(g, n, active=full, normalize=True, variant="hybrid"))
timed("allreduce twoshot 8 x 16 MiB", lambda: ce.allreduce(g, n, active=full, normalize=True, variant="twoshot"))
cur = [0]


def ea():
    a, b = (c0, c1) if cur[0] == 0 else (c1, c0)
    ce.fused_ea(p, a, b, n, 0.1, active=full, variant="hybrid")
    cur[0] ^= 1


timed("fused_ea hybrid (all P2P) 8 x 16 MiB", ea)
timed("broadcast wide two-shot 8 x 16 MiB", lambda: ce.broadcast(p, n, root=3, variant="twoshot"))
ns = ce.ll_max_bytes // 4
timed(f"fused_sgd LL 8 x {ns * 4 >> 10} KiB", lambda: ce.fused_sgd(g, p, None, ns, 1e-3, active=full, zero_grads=True, variant="ll"))
timed(f"allreduce LL 8 x {ns * 4 >> 10} KiB", lambda: ce.allreduce(g, ns, active=full, normalize=True, variant="ll"))

# ---- asynchronous EASGD client kernel: this GPU is server and client at once --------------------------------------
ca = dl.EmulatedCommunicator(2, dev, heap_mb=40)
ca._c.emulate = 0
ca._c.rank = 1
center, _ = ca.alloc_all(n)
xa = torch.randn(n, device=dev)
timed("async_ea_client 16 MiB (lock-free)", lambda: ca.async_client_sync(0, xa, center, n, 0.2, exclusive=False))

# ---- tail kernels of the flagship model, layer-2 shape: [128, 16, 16, 128] -----------------------------------------
B, H, C = 128, 16, 128
y = torch.randn(B, H, H, C, device=dev)
dy = torch.empty_like(y)
dp = torch.randn(B, H // 2, H // 2, C, device=dev)
pl = torch.empty_like(dp)
gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
mean, invstd = torch.empty(C, device=dev), torch.empty(C, device=dev)
dgamma, dbeta = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
work = torch.zeros(int(lib.dl_bn_work_bytes(C)) // 8 + 1, dtype=torch.float64, device=dev)
st = lambda: torch.cuda.current_stream().cuda_stream  # noqa: E731
timed("bn+relu+pool forward  [128,16,16,128]", lambda: _lib.check(lib.dl_bn_relu_pool_fwd(
    y.data_ptr(), pl.data_ptr(), gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(), invstd.data_ptr(), None, None, B, H, H, C, 1e-3, 0.1,
    work.data_ptr(), st())))
timed("bn+relu+pool backward [128,16,16,128]", lambda: _lib.check(lib.dl_bn_relu_pool_bwd(
    y.data_ptr(), dp.data_ptr(), dy.data_ptr(), gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(), invstd.data_ptr(), dgamma.data_ptr(),
    dbeta.data_ptr(), B, H, H, C, 1, work.data_ptr(), st())))
timed("colsum [32768, 128]", lambda: _lib.check(lib.dl_colsum(dy.data_ptr(), dbeta.data_ptr(), B * H * H, C, 1, work.data_ptr(), st())))
x3 = torch.randn(B, 32, 32, 3, device=dev)
col = torch.empty(B * 32 * 32, 76, device=dev)
timed("im2col5x5 [128,32,32,3]", lambda: _lib.check(lib.dl_im2col5x5(x3.data_ptr(), col.data_ptr(), B, 32, 32, 3, 76, st())))
from torch_distlearn_b200.ops.head import head_nll  # noqa: E402
feat = torch.randn(B, 2, 2, 512, device=dev, requires_grad=True)
hw = torch.nn.Parameter(torch.randn(10, 2048, device=dev) * 0.02)
hb = torch.nn.Parameter(torch.zeros(10, device=dev))
hw.grad, hb.grad = torch.zeros_like(hw), torch.zeros_like(hb)
tgt = torch.randint(0, 10, (B,), device=dev)


def head():
    loss, _ = head_nll(feat, hw, hb, tgt)
    loss.backward()


timed("head forward+backward [128, 2048] -> 10", head)
print("done", flush=True)
