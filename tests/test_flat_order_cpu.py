"""FlatParams(order=...): the layout of the flat vector is a choice (module order like getParameters(), execution order
for layer-by-layer buckets, or an explicit list).  Training must not depend on it, checkpoints must move between
layouts by name, and in execution order every bucket of GradBuckets must be ready as soon as ITS layers are done."""
import os
import uuid

import pytest
import torch
import torch.nn.functional as F

import torch_distlearn_b200 as dl
from torch_distlearn_b200.parallel.bucketed import GradBuckets


class _Net(torch.nn.Module):
    """Registers the 'convolutions' first and the 'norm' scales after, like the CIFAR model does."""

    def __init__(self, seed=0):
        super().__init__()
        torch.manual_seed(seed)
        self.lins = torch.nn.ModuleList([torch.nn.Linear(12, 12) for _ in range(3)])
        self.scales = torch.nn.ParameterList([torch.nn.Parameter(torch.rand(12) + 0.5) for _ in range(3)])
        self.head = torch.nn.Linear(12, 4)

    def execution_order(self):
        names = []
        for i in range(3):
            names += [f"lins.{i}.weight", f"lins.{i}.bias", f"scales.{i}"]
        return names + ["head.weight", "head.bias"]

    def forward(self, x):
        for lin, s in zip(self.lins, self.scales):
            x = torch.tanh(lin(x) * s)
        return F.log_softmax(self.head(x), dim=1)


def _comm():
    return dl.HostCommunicator(0, 1, name=f"fo{os.getpid()}{uuid.uuid4().hex[:6]}", heap_mb=8, timeout_s=30)


def _train(order, bucket_mb=0.0, steps=4):
    comm = _comm()
    try:
        model = _Net()
        flat = dl.FlatParams(comm, model, with_momentum=True, order=order)
        sgd = dl.AllReduceSGD(comm, flat)
        buckets = None
        if bucket_mb:
            flat.attach_grads()
            buckets = GradBuckets(flat, sgd, bucket_mb=bucket_mb)
            buckets.set_hyper(lr=0.1, momentum=0.9, weight_decay=1e-3)
        g = torch.Generator().manual_seed(1)
        ready_at = []
        for _ in range(steps):
            x, y = torch.randn(6, 12, generator=g), torch.randint(0, 4, (6,), generator=g)
            F.nll_loss(model(x), y).backward()
            if buckets is not None:
                ready_at.append(buckets.launched_from_hooks)
                buckets.finish()
            else:
                sgd.step(lr=0.1, momentum=0.9, weight_decay=1e-3)
        out = {n: p.detach().clone() for n, p in model.named_parameters()}
        info = dict(names=list(flat.names), nb=len(buckets.buckets) if buckets else 0, ready_at=ready_at,
                    members=[b.n_params for b in buckets.buckets] if buckets else [])
        if buckets is not None:
            buckets.remove()
        return out, info
    finally:
        comm.close()


def test_layout_is_what_was_asked_for():
    comm = _comm()
    try:
        model = _Net()
        flat = dl.FlatParams(comm, model, order="execution")
        assert flat.names == model.execution_order()
        assert flat.offsets == sorted(flat.offsets) and flat.offsets[0] == 0
        for n, p in model.named_parameters():                              # every parameter is a view at its slot
            i = flat.names.index(n)
            assert p.data_ptr() == flat.params.tensor.data_ptr() + 4 * flat.offsets[i]
            assert p.grad.data_ptr() == flat.grads.tensor.data_ptr() + 4 * flat.offsets[i]
        with pytest.raises(ValueError):
            dl.FlatParams(comm, _Net(), order=["head.weight"])
        with pytest.raises(ValueError):
            dl.FlatParams(comm, _Net(), order="alphabetical")
        explicit = list(reversed(model.execution_order()))
        assert dl.FlatParams(comm, _Net(), order=explicit).names == explicit
        assert dl.FlatParams(comm, torch.nn.Linear(3, 3), order="execution").names == ["weight", "bias"]   # no hint: module order
    finally:
        comm.close()


def test_training_does_not_depend_on_the_layout():
    base, _ = _train("module")
    for order, bucket_mb in (("execution", 0.0), ("module", 2e-4), ("execution", 2e-4)):
        got, info = _train(order, bucket_mb)
        for n in base:
            assert torch.equal(got[n], base[n]), (order, bucket_mb, n)


def test_execution_order_lets_buckets_leave_during_backward():
    _, mod = _train("module", 6e-4, steps=1)
    _, exe = _train("execution", 6e-4, steps=1)
    assert mod["nb"] >= 3 and exe["nb"] >= 3
    # in execution order every bucket is complete when backward has passed its layers: all leave from hooks
    assert exe["ready_at"][0] == exe["nb"]
    # in module order the first bucket (head + all the scales) waits for layer 0's scale, i.e. almost all of backward;
    # it still leaves from a hook, but nothing can overlap it — what is asserted here is only that the two runs agree
    assert mod["ready_at"][0] >= mod["nb"] - 1


def test_checkpoint_moves_between_layouts(tmp_path):
    comm = _comm()
    try:
        a, b = _Net(seed=1), _Net(seed=2)
        fa = dl.FlatParams(comm, a, with_momentum=True, name="a", order="module")
        fb = dl.FlatParams(comm, b, with_momentum=True, name="b", order="execution")
        fa.momentum.tensor.copy_(fa.params.tensor * 3)
        path = str(tmp_path / "layout.dlb200")
        dl.save_checkpoint(path, fa, momentum=fa.momentum)
        dl.load_checkpoint(path, fb, momentum=fb.momentum)
        for (n, p), (_, q) in zip(a.named_parameters(), b.named_parameters()):
            assert torch.equal(p, q), n
        for n in fa.names:
            ia, ib = fa.names.index(n), fb.names.index(n)
            za = fa.sizes[ia]
            assert torch.equal(fb.momentum.tensor[fb.offsets[ib]:fb.offsets[ib] + za],
                               fa.momentum.tensor[fa.offsets[ia]:fa.offsets[ia] + za])
    finally:
        comm.close()
