"""GradBuckets bookkeeping on the CPU transport (world 1, in-process): a backward function that writes ``.grad``
itself and reports it with ``grad_done`` must make its bucket leave exactly once per step, whether or not torch also
runs the post-accumulate hooks for the ``None`` gradient it returns (torch 2.11 does)."""
import os
import uuid

import torch

import torch_distlearn_b200 as dl
from torch_distlearn_b200.ops.bn import grad_done
from torch_distlearn_b200.parallel.bucketed import GradBuckets


class _ScaleDirect(torch.autograd.Function):
    """y = x * w with the weight gradient added into w.grad by the backward itself (what the CUDA backward kernels do)."""

    @staticmethod
    def forward(ctx, x, w):
        ctx.save_for_backward(x, w)
        ctx.param = w
        return x * w

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        ctx.param.grad.add_((gy * x).sum(0))
        grad_done(ctx.param)
        return gy * w, None


class _Net(torch.nn.Module):
    def __init__(self):
        super().__init__()
        torch.manual_seed(3)
        self.a = torch.nn.Parameter(torch.randn(8))
        self.lin = torch.nn.Linear(8, 8)
        self.b = torch.nn.Parameter(torch.randn(8))

    def forward(self, x):
        return _ScaleDirect.apply(self.lin(_ScaleDirect.apply(x, self.a)), self.b).sum()


def test_direct_grad_writers_launch_each_bucket_once_per_step():
    comm = dl.HostCommunicator(0, 1, name=f"bk{os.getpid()}{uuid.uuid4().hex[:6]}", heap_mb=8, timeout_s=30)
    try:
        model, ref = _Net(), _Net()
        flat = dl.FlatParams(comm, model, with_momentum=True)
        sgd = dl.AllReduceSGD(comm, flat)
        flat.attach_grads()
        # ~20-element buckets, the first one holds b and at least one of lin's tensors.  A second count for b (torch's hook for the None gradient)
        # would send that bucket off before lin's gradient exists
        buckets = GradBuckets(flat, sgd, bucket_mb=20 * 4 / 2**20)
        buckets.set_hyper(lr=0.1, momentum=0.9)
        assert len(buckets.buckets) >= 2 and buckets.buckets[0].n_params >= 2, [(b.start, b.numel, b.n_params) for b in buckets.buckets]
        opt = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9)
        g = torch.Generator().manual_seed(0)
        for step in range(3):
            x = torch.randn(4, 8, generator=g)
            before = comm.launches
            model(x).backward()
            buckets.finish()
            assert comm.launches - before == len(buckets.buckets)        # not one more, not one fewer
            assert float(flat.grads.tensor.abs().max()) == 0.0
            for p in ref.parameters():
                p.grad = torch.zeros_like(p)
            ref(x).backward()
            opt.step()
            for p, q in zip(model.parameters(), ref.parameters()):
                torch.testing.assert_close(p.detach(), q.detach(), rtol=1e-5, atol=1e-6)
        # every bucket but (at most) the last left from a hook during backward
        assert buckets.launched_from_hooks >= 3 * (len(buckets.buckets) - 1)
        assert sgd.steps == 3
        buckets.remove()
    finally:
        comm.close()
