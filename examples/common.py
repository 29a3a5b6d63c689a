"""Helpers shared by the example programs (the reference shares examples/Model.lua and
examples/Data.lua between its scripts; the models and datasets live in the package here)."""
from __future__ import annotations

import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch
import torch.nn.functional as F


def setup(opt, rank=None, world=None, loud=False):
    """Communicator + rank-0 logger from the reference-style flags."""
    from torch_distlearn_b200.launch import make_communicator
    from torch_distlearn_b200.utils import RankLogger
    torch.manual_seed(opt.seed)
    torch.set_num_threads(int(os.environ.get("OMP_NUM_THREADS", "1")))
    comm = make_communicator(opt, rank=rank, world=world)
    # nodes > 1 stay silent (examples/mnist.lua:20-23) unless the role wants its own voice
    log = RankLogger(1 if loud else comm.node_index, opt.metrics or None)
    return comm, log


def easgd_topology(opt):
    """Ranks of an asynchronous EASGD job with N = --numNodes clients (examples/AsyncEASGD.sh:36-40):
    rank 0 = server, ranks 1..N = clients (client i has --nodeIndex i), rank N+1 = tester (optional)."""
    world = opt.numNodes + 1 + (1 if opt.tester else 0)
    return world, 0, (opt.numNodes + 1 if opt.tester else None)


def min_over_nodes(comm, value: int) -> int:
    """Smallest of an integer over all nodes, through the barrier payload (no extra collective)."""
    if comm.world == 1:
        return value
    res = comm.barrier(active=1, payload=min(int(value), 2**30 - 1), channel=1, read=True)
    return min(res.payloads)


def batches_this_epoch(num_batches: int, opt) -> int:
    return min(num_batches, opt.epochSize) if opt.epochSize > 0 else num_batches


def forward_backward(model, batch):
    out = model(batch.input)
    loss = F.nll_loss(out, batch.target)        # ClassNLLCriterion on log-probabilities
    loss.backward()
    return out.detach(), loss.detach()


class MnistKernelStep:
    """forward + backward of the MNIST network as ONE kernel (csrc/mnist_step.cu, an 8-CTA cluster) instead of ~25 small
    library kernels; with ``lr`` given (and a batch of at most 8) the plain SGD update is fused in as well.  Returns
    what ``forward_backward`` returns: (log-probabilities, loss), both on the device."""

    def __init__(self, model, batch_size: int):
        from torch_distlearn_b200.ops.mnist import model_tensors
        self.params, self.grads = model_tensors(model)
        dev = self.params[0].device
        self.loss = torch.zeros(1, device=dev)
        self.logp = torch.zeros(batch_size, 10, device=dev)

    @staticmethod
    def usable(model, device, batch_size: int) -> bool:
        return device.type == "cuda" and type(model).__name__ == "MnistConvNet" and batch_size <= 4096

    def __call__(self, batch, lr=None):
        from torch_distlearn_b200.ops.mnist import FUSED_MAX_BATCH, mnist_step
        B = batch.input.shape[0]
        fuse = lr is not None and B <= FUSED_MAX_BATCH
        x = batch.input.reshape(B, 1024).contiguous()
        mnist_step(x, batch.target.to(torch.int64).contiguous(), self.params, self.grads, self.loss, self.logp[:B],
                   lr=lr or 0.0, apply_sgd=fuse)
        return self.logp[:B], self.loss[0], fuse


@torch.no_grad()
def evaluate(model, get_batch, n_batches: int, confusion, batch_stats: bool = False) -> float:
    """Mean loss over ``n_batches``.  ``batch_stats`` keeps BatchNorm on batch statistics: a model that
    only ever received parameters (the EASGD tester pulls the center, and BatchNorm running statistics
    are not parameters — SURVEY §2.7) has no meaningful running statistics to evaluate with."""
    was_training = model.training
    model.train(batch_stats)
    total, n = 0.0, 0
    for _ in range(n_batches):
        b = get_batch()
        out = model(b.input)
        total += float(F.nll_loss(out, b.target, reduction="sum"))
        n += b.batchSize
        confusion.batchAdd(out, b.target)
    model.train(was_training)
    return total / max(n, 1)


class Throughput:
    def __init__(self):
        self.t0, self.samples = time.time(), 0

    def add(self, n: int) -> None:
        self.samples += n

    def rate(self) -> float:
        return self.samples / max(time.time() - self.t0, 1e-9)
