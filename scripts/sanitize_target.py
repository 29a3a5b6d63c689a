"""Tiny single-GPU workload touching every kernel family, for `compute-sanitizer --tool memcheck` (one tool per call)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch_distlearn_b200 as dl
from torch_distlearn_b200.ops import FusedBNReLUPool, TCConv5x5, tc_gemm, sumsq_diff, local_sgd

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
a = torch.randn(256, 128, device=dev); b = torch.randn(128, 128, device=dev)
c = tc_gemm(a, b)
c2 = tc_gemm(a.T.contiguous(), b.T.contiguous(), a_mn=True, b_mn=True, split_k=2)
print("gemm", float((c - a @ b.T).abs().max()), float((c2 - a @ b.T).abs().max()))
for cin, cout, H in [(3, 32, 8), (32, 32, 8)]:
    conv = TCConv5x5(cin, cout).to(dev)
    bn = FusedBNReLUPool(cout).to(dev)
    x = torch.randn(2, H, H, cin, device=dev, requires_grad=True)
    bn(conv(x)).sum().backward()
    print("conv/bn", cin, cout, float(x.grad.abs().sum()))
comm = dl.EmulatedCommunicator(2, dev, heap_mb=4)
g, gv = comm.alloc_all(4096); p, pv = comm.alloc_all(4096); m, mv = comm.alloc_all(4096)
c0, _ = comm.alloc_all(4096); c1, _ = comm.alloc_all(4096)
for v in gv:
    v.normal_()
for variant in ("oneshot", "twoshot"):
    comm.fused_sgd(g, p, m, 4096, 0.1, momentum=0.9, active=0b11, variant=variant)
    comm.allreduce(g, 4096, active=0b11, normalize=True, variant=variant)
    comm.broadcast(p, 4096, root=1, also=c0, variant=variant)
comm.fused_ea(p, c0, c1, 4096, 0.1, active=0b11, variant="twoshot")
torch.cuda.synchronize()
print("collectives", float(pv[0].abs().sum()), comm.result_of(0).n_active)
x = torch.randn(10007, device=dev); y = torch.randn(10007, device=dev)
local_sgd(x, y, None, lr=0.1)
print("local", float(sumsq_diff(x, y)))
print("SANITIZE TARGET OK")
