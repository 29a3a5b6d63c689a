"""torch.distributed baselines of AllReduceSGD / AllReduceEA / AsyncEA.

They mirror the reference's dataflow with library collectives:

* AllReduceSGD : all_reduce(flat grads) -> div_(n) -> torch SGD           (lua/AllReduceSGD.lua:18-30)
* AllReduceEA  : delta = a(x-c); x -= delta; all_reduce(delta); c += delta  (lua/AllReduceEA.lua:34-45)
* AsyncEA      : a server rank runs the host loop recv(any) -> send center -> recv delta -> add
                 (lua/AsyncEA.lua:230-237) over point-to-point send/recv.
"""
from __future__ import annotations

from typing import Iterable, Optional

import torch
import torch.distributed as dist


def _flatten_like(params: Iterable[torch.Tensor]):
    params = list(params)
    total = sum(p.numel() for p in params)
    flat = torch.zeros(total, dtype=params[0].dtype, device=params[0].device)
    views, off = [], 0
    for p in params:
        n = p.numel()
        views.append(flat[off:off + n].view(p.shape))
        off += n
    return flat, views


class NcclFlat:
    """Flat parameter / gradient vectors (plain device memory), params and grads are views."""

    def __init__(self, model: torch.nn.Module):
        self.model_params = list(model.parameters())
        self.params, pviews = _flatten_like(self.model_params)
        self.grads, gviews = _flatten_like(self.model_params)
        with torch.no_grad():
            for p, pv, gv in zip(self.model_params, pviews, gviews):
                pv.copy_(p)
                p.data = pv
                p.grad = gv


class NcclAllReduceSGD:
    def __init__(self, flat: NcclFlat, group=None):
        self.flat, self.group = flat, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1

    def synchronize_parameters(self):
        if self.world > 1:
            dist.broadcast(self.flat.params, src=0, group=self.group)

    def step(self, lr: float, momentum: float = 0.0, momentum_buf: Optional[torch.Tensor] = None):
        g, p = self.flat.grads, self.flat.params
        if self.world > 1:
            dist.all_reduce(g, op=dist.ReduceOp.AVG, group=self.group)   # NCCL divides in the reduction
        if momentum and momentum_buf is not None:
            momentum_buf.mul_(momentum).add_(g)
            p.add_(momentum_buf, alpha=-lr)
        else:
            p.add_(g, alpha=-lr)
        g.zero_()


class NcclAllReduceEA:
    def __init__(self, flat: NcclFlat, tau: int, alpha: float, group=None):
        self.flat, self.tau, self.alpha, self.group = flat, tau, alpha, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.center = flat.params.clone()
        self.delta = torch.zeros_like(flat.params)
        self.step_count = 0

    def average_parameters(self) -> bool:
        self.step_count += 1
        if self.step_count % self.tau:
            return False
        x = self.flat.params
        torch.sub(x, self.center, out=self.delta)
        self.delta.mul_(self.alpha)
        x.sub_(self.delta)
        if self.world > 1:
            dist.all_reduce(self.delta, group=self.group)
        self.center.add_(self.delta)
        return True

    def synchronize_center(self):
        if self.world > 1:
            dist.broadcast(self.center, src=0, group=self.group)


class HostLoopAsyncEA:
    """Reference-shaped async EASGD over torch.distributed point-to-point: rank 0 is the server and
    serves one client at a time (recv from any source = the mutex)."""

    def __init__(self, flat_params: torch.Tensor, tau: int, alpha: float, server: int = 0):
        self.x, self.tau, self.alpha, self.server = flat_params, tau, alpha, server
        self.rank = dist.get_rank()
        self.center = flat_params.clone()
        self.delta = torch.zeros_like(flat_params)
        self.req = torch.zeros(1, dtype=torch.int64, device=flat_params.device)
        self.step_count = 0

    def sync_client(self) -> bool:
        self.step_count += 1
        if self.step_count % self.tau:
            return False
        self.req.fill_(self.rank)
        dist.send(self.req, dst=self.server)              # "Enter?"
        dist.recv(self.center, src=self.server)           # center
        torch.sub(self.x, self.center, out=self.delta)
        self.delta.mul_(self.alpha)
        self.x.sub_(self.delta)
        dist.send(self.delta, dst=self.server)            # delta
        return True

    def sync_server(self) -> int:
        src = dist.recv(self.req)                         # recvAny
        dist.send(self.center, dst=src)
        dist.recv(self.delta, src=src)
        self.center.add_(self.delta)
        return src
