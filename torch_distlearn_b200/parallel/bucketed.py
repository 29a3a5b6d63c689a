"""GradBuckets — the fused all-reduce + SGD kernel launched per bucket, overlapped with backward.

The reference waits for the whole backward pass, then all-reduces every gradient tensor one after
another (examples/cifar10.lua:181-191).  Here the flat gradient vector is cut into contiguous
buckets in *backward order* (last layer first).  As soon as autograd has accumulated the last
gradient of a bucket, a post-accumulate hook records an event on the compute stream and launches
``fused_allreduce_normalize_sgd`` for that slice of the flat vector on a side stream, with a capped
number of CTAs: communication and the parameter update of layer L run under the backward kernels of layers < L.
On the CUDA backends the launches use the LIGHT variant of the kernel (128 threads, <= 80 registers per CTA): small
enough to sit on an SM next to the tensor-core GEMM's persistent CTA, so that the bucket really runs beside the
convolutions instead of queueing for whole SMs (csrc/dl_pipe.cuh); buckets below the small-message limit take the
push path (one NVLink crossing, no barrier).  Buckets are always launched in the same fixed order on every
rank (a bucket that becomes ready early waits for its predecessors), because every rank must issue
the same sequence of collective kernels.  The whole fork/join pattern is capturable in a CUDA graph.
"""
from __future__ import annotations

from typing import List

import torch

from .allreduce_sgd import AllReduceSGD
from .flat import FlatParams


class _Bucket:
    __slots__ = ("start", "numel", "n_params", "pending", "grads", "params", "momentum")


class GradBuckets:
    def __init__(self, flat: FlatParams, sgd: AllReduceSGD, bucket_mb: float = 8.0, nblocks: int = 0, channel: int = 2,
                 variant: str = "auto"):
        if flat.grads is None:
            raise ValueError("GradBuckets needs FlatParams with gradients")
        self.flat, self.sgd, self.comm = flat, sgd, sgd.comm
        self.nblocks, self.channel, self.variant = nblocks, channel, variant
        self.device = flat.device
        self.stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self._event = torch.cuda.Event() if self.device.type == "cuda" else None
        self.buckets: List[_Bucket] = []
        self._bucket_of_param: List[int] = [0] * len(flat.sizes)
        limit = max(int(bucket_mb * (1 << 20)) // 4, 4)
        # walk the parameters in backward order; a bucket is the contiguous range they cover
        end = flat.padded
        count, members = 0, []
        order = list(range(len(flat.sizes) - 1, -1, -1))
        for k, i in enumerate(order):
            members.append(i)
            count += flat.sizes[i]
            last = k == len(order) - 1
            if count >= limit or last:
                start = 0 if last else flat.offsets[i]
                b = _Bucket()
                b.start, b.numel, b.n_params, b.pending = start, end - start, len(members), len(members)
                b.grads = flat.grads.sub(start, b.numel)
                b.params = flat.params.sub(start, b.numel)
                b.momentum = flat.momentum.sub(start, b.numel) if flat.momentum is not None else None
                for m in members:
                    self._bucket_of_param[m] = len(self.buckets)
                self.buckets.append(b)
                end, count, members = start, 0, []
        self._next = 0                      # next bucket to launch (fixed order)
        self._fired = [False] * len(flat.sizes)
        self.launched_from_hooks = 0        # buckets that went out DURING backward (the overlap), for tests / diagnostics
        self._hyper = dict(lr=0.01, momentum=0.0, weight_decay=0.0, nesterov=False)
        self._handles = []
        for i, p in enumerate(flat._params):
            if p.requires_grad:
                self._handles.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))
            else:
                self.buckets[self._bucket_of_param[i]].pending -= 1
        for b in self.buckets:
            b.n_params = b.pending
        self.enabled = True
        # a rank whose partition is empty must still serve its peers' per-bucket launches when draining
        sgd._last = ("bucketed", self, None, None)

    def _make_hook(self, index: int):
        def hook(_param):
            # once per parameter per step: the hook can be reached both from a backward kernel that wrote .grad itself
            # (ops.bn.grad_done) and from autograd's own accumulation node
            if not self.enabled or self._fired[index]:
                return
            self._fired[index] = True
            b = self.buckets[self._bucket_of_param[index]]
            b.pending -= 1
            before = self._next
            self._launch_ready()
            self.launched_from_hooks += self._next - before
        return hook

    def set_hyper(self, lr: float, momentum: float = 0.0, weight_decay: float = 0.0, nesterov: bool = False) -> None:
        if momentum != 0.0 and self.flat.momentum is None:
            raise ValueError("momentum needs FlatParams(with_momentum=True)")
        self._hyper = dict(lr=lr, momentum=momentum, weight_decay=weight_decay, nesterov=nesterov)

    # ------------------------------------------------------------------------------------------------
    def _launch(self, b: _Bucket, active: int) -> None:
        h = self._hyper
        payload = min(self.sgd.steps + (1 if active else 0), 2**30 - 1)
        variant = self.variant
        if variant == "auto" and self.stream is not None and b.numel * 4 > getattr(self.comm, "ll_max_bytes", 0):
            variant = "light"            # shares the SMs with the backward kernels it overlaps
        kw = dict(momentum=h["momentum"], weight_decay=h["weight_decay"], nesterov=h["nesterov"], active=active,
                  payload=payload, zero_grads=True, variant=variant, channel=self.channel, read=False)
        mom = b.momentum if h["momentum"] != 0.0 else None
        if self.stream is None:
            self.comm.fused_sgd(b.grads, b.params, mom, b.numel, h["lr"], **kw)
            return
        self._event.record(torch.cuda.current_stream(self.device))
        self.stream.wait_event(self._event)
        self.comm.fused_sgd(b.grads, b.params, mom, b.numel, h["lr"], nblocks=self.nblocks, stream=self.stream, **kw)

    def _launch_ready(self) -> None:
        while self._next < len(self.buckets) and self.buckets[self._next].pending <= 0:
            self._launch(self.buckets[self._next], active=1)
            self._next += 1

    def finish(self) -> None:
        """Call after ``backward()``: launches whatever is left (parameters that received no gradient
        contribute their zeros), joins the side stream and counts the step."""
        while self._next < len(self.buckets):
            self._launch(self.buckets[self._next], active=1)
            self._next += 1
        if self.stream is not None:
            torch.cuda.current_stream(self.device).wait_stream(self.stream)
        self._next = 0
        self._fired = [False] * len(self._fired)
        for b in self.buckets:
            b.pending = b.n_params
        self.sgd.steps += 1
        self.sgd._last = ("bucketed", self, None, None)

    def drain_round(self):
        """One round of ``active = 0`` launches with this geometry (a finished rank serving its
        peers, lua/AllReduceSGD.lua:37); returns the result of the last bucket's barrier."""
        for b in self.buckets:
            self._launch(b, active=0)
        if self.stream is not None:
            torch.cuda.current_stream(self.device).wait_stream(self.stream)
        return self.comm.read_result(self.channel, self.stream) if self.stream is not None else self.comm.read_result(self.channel)

    def remove(self) -> None:
        for h in self._handles:
            h.remove()
        self._handles = []
