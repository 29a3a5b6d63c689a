"""Diagnostic: per-parameter difference between the whole-vector fused step and the per-bucket (GradBuckets) step of
the tensor-core CIFAR model, under a few configurations (direct .grad accumulation on/off, side stream serialized).
Run on one GPU: python scripts/bucket_probe.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import torch_distlearn_b200 as dl
from torch_distlearn_b200.models import cifar_convnet_tc
from torch_distlearn_b200.ops import bn as bn_ops
from torch_distlearn_b200.parallel.trainer import TrainStep

DEV = "cuda:0"


def run(bucket_mb, direct=True, serialize=False, steps=4, trace=None):
    bn_ops.DIRECT_GRAD = direct
    torch.manual_seed(0)
    comm = dl.P2PCommunicator(rank=0, world=1, device=torch.device(DEV), heap_mb=96)
    model = cifar_convnet_tc().to(DEV)
    flat = dl.FlatParams(comm, model, with_momentum=True)
    sgd = dl.AllReduceSGD(comm, flat)
    step = TrainStep(model, flat, sgd, (16, 3, 32, 32), lr=0.002, momentum=0.9, use_graph=False, warmup=1,
                     bucket_mb=bucket_mb, bucket_blocks=8)
    if serialize and step.buckets is not None:
        inner = step.buckets._launch

        def launch(b, active):
            inner(b, active)
            torch.cuda.current_stream().wait_stream(step.buckets.stream)
        step.buckets._launch = launch
    g = torch.Generator().manual_seed(5)
    if os.environ.get("PROBE_ZERO_WARMUP", "0") != "1":        # default: warm up on a real batch, not on the zero buffers
        step.load(torch.randn(16, 3, 32, 32, generator=g), torch.randint(0, 10, (16,), generator=g))
    step.capture()
    snaps = [flat.params.tensor.clone()]
    for _ in range(steps):
        step.step(torch.randn(16, 3, 32, 32, generator=g).to(DEV), torch.randint(0, 10, (16,), generator=g).to(DEV))
        torch.cuda.synchronize()
        snaps.append(flat.params.tensor.clone())
    dirty = [n for n, gv in zip(flat.names, flat.grad_views) if float(gv.abs().max()) != 0.0]
    info = dict(names=list(flat.names), offsets=list(flat.offsets), sizes=list(flat.sizes), dirty=dirty,
                hooks=step.buckets.launched_from_hooks if step.buckets else None,
                nb=len(step.buckets.buckets) if step.buckets else 0)
    if step.buckets is not None:
        step.buckets.remove()
    comm.close() if hasattr(comm, "close") else None
    return snaps, info


def compare(tag, base, other, info):
    print(f"== {tag}: buckets={info['nb']} launched_from_hooks={info['hooks']} dirty={info['dirty']}", flush=True)
    for s, (a, b) in enumerate(zip(base, other)):
        worst = []
        for n, o, sz in zip(info["names"], info["offsets"], info["sizes"]):
            d = float((a[o:o + sz] - b[o:o + sz]).abs().max())
            if d > 1e-5:
                worst.append(f"{n}:{d:.2e}")
        print(f"   after step {s}: max {float((a - b).abs().max()):.3e}  {' '.join(worst[:8])}", flush=True)


base, info0 = run(0.0)
again, info = run(0.0)
compare("whole-vector run repeated (determinism of the model itself)", base, again, info)
for tag, kw in [("buckets, direct .grad", dict(direct=True)),
                ("buckets, autograd accumulation", dict(direct=False)),
                ("buckets, direct .grad, side stream serialized", dict(direct=True, serialize=True))]:
    snaps, info = run(2.0, **kw)
    compare(tag, base, snaps, info)
base_nd, info = run(0.0, direct=False)
compare("whole-vector, autograd accumulation vs direct", base, base_nd, info)
