#!/usr/bin/env python
"""ncu target: the one-launch MNIST step and the tail kernels of the flagship model that the first capture did not reach
(BatchNorm backward, column sums, classifier head, im2col), a few launches each.

    python scripts/ncu_tails_target.py && ncu --set full --clock-control none --import-source on \
        -k regex:"mnist_step|bn_relu_pool_bwd|bn_bwd_finalize|colsum|head_|im2col" -c 30 -o gpurun_out/prof_tails -f python scripts/ncu_tails_target.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from torch_distlearn_b200 import _lib
from torch_distlearn_b200.models import mnist_convnet
from torch_distlearn_b200.ops.head import head_nll
from torch_distlearn_b200.ops.mnist import mnist_step, model_tensors

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
lib = _lib.cuda()
st = lambda: torch.cuda.current_stream().cuda_stream  # noqa: E731


def timed(name, fn, iters=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:52s} {e0.elapsed_time(e1) / iters * 1e3:9.1f} us", flush=True)


model = mnist_convnet().to(dev)
for p in model.parameters():
    p.grad = torch.zeros_like(p)
params, grads = model_tensors(model)
loss = torch.zeros(1, device=dev)
for B in (1, 8):
    x = torch.randn(B, 1024, device=dev)
    y = torch.randint(0, 10, (B,), device=dev)
    timed(f"mnist step, batch {B} (forward+backward+SGD)", lambda: mnist_step(x, y, params, grads, loss, None, lr=0.01, apply_sgd=True))

B, H, C = 128, 16, 128
yv = torch.randn(B, H, H, C, device=dev)
dy = torch.empty_like(yv)
dp = torch.randn(B, H // 2, H // 2, C, device=dev)
gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
mean, invstd = torch.zeros(C, device=dev), torch.ones(C, device=dev)
dgamma, dbeta, dcb = torch.zeros(C, device=dev), torch.zeros(C, device=dev), torch.zeros(C, device=dev)
work = torch.zeros(int(lib.dl_bn_work_bytes(C)) // 8 + 1, dtype=torch.float64, device=dev)
timed("bn+relu+pool backward + conv-bias gradient", lambda: _lib.check(lib.dl_bn_relu_pool_bwd_bias(
    yv.data_ptr(), dp.data_ptr(), dy.data_ptr(), gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(), invstd.data_ptr(), dgamma.data_ptr(),
    dbeta.data_ptr(), dcb.data_ptr(), B, H, H, C, 1, work.data_ptr(), st())))
timed("colsum [32768, 128]", lambda: _lib.check(lib.dl_colsum(dy.data_ptr(), dbeta.data_ptr(), B * H * H, C, 1, work.data_ptr(), st())))
x3 = torch.randn(B, 32, 32, 3, device=dev)
col = torch.empty(B * 32 * 32, 76, device=dev)
timed("im2col5x5 [128,32,32,3]", lambda: _lib.check(lib.dl_im2col5x5(x3.data_ptr(), col.data_ptr(), B, 32, 32, 3, 76, st())))
feat = torch.randn(B, 2, 2, 512, device=dev, requires_grad=True)
hw = torch.nn.Parameter(torch.randn(10, 2048, device=dev) * 0.02)
hb = torch.nn.Parameter(torch.zeros(10, device=dev))
hw.grad, hb.grad = torch.zeros_like(hw), torch.zeros_like(hb)
tgt = torch.randint(0, 10, (B,), device=dev)


def head():
    l, _ = head_nll(feat, hw, hb, tgt)
    l.backward()


timed("head forward+backward", head)
print("done", flush=True)
