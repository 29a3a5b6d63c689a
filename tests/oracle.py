"""Single-process N-node oracle: a CPU restatement of the reference's math.

Every kernel / transport test compares against these functions.  They mirror the *semantics* of
lua/AllReduceSGD.lua, lua/AllReduceEA.lua and lua/AsyncEA.lua on plain fp32 torch tensors, using
the same evaluation order as the kernels (sum over ranks in rank order, then scale) so that
results can be compared tightly.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch


def allreduce_sum(values: Sequence[torch.Tensor], active: Optional[Sequence[int]] = None, normalize: bool = False):
    """tree.allReduce(value, add): returns (result, n_active).  Inactive nodes contribute zeros
    (lua/AllReduceSGD.lua:37); active==2 contributes but is not counted."""
    n = len(values)
    active = list(active) if active is not None else [1] * n
    acc = torch.zeros_like(values[0])
    for v, a in zip(values, active):
        if a:
            acc = acc + v
    n_active = sum(1 for a in active if a & 1)
    if normalize and n_active > 1:
        acc = acc * torch.tensor(1.0 / n_active, dtype=acc.dtype) if acc.is_floating_point() else acc
    return acc, n_active


def fused_sgd(grads: Sequence[torch.Tensor], params: torch.Tensor, lr: float, active: Optional[Sequence[int]] = None,
              momentum: float = 0.0, momentum_buf: Optional[torch.Tensor] = None, weight_decay: float = 0.0,
              nesterov: bool = False):
    """sumAndNormalizeGradients + p:add(-lr, g) (lua/AllReduceSGD.lua:18-30; examples/cifar10.lua:187-191).
    Returns (new_params, new_momentum, n_active); params unchanged when nobody is active."""
    g, n_active = allreduce_sum(grads, active, normalize=True)
    if n_active == 0:
        return params.clone(), momentum_buf, 0
    p = params.clone()
    if weight_decay != 0.0:
        g = g + weight_decay * p
    m = momentum_buf
    if momentum_buf is not None:
        m = momentum * momentum_buf + g
        g = g + momentum * m if nesterov else m
    p = p - lr * g
    return p, m, n_active


def elastic_average(xs: Sequence[torch.Tensor], center: torch.Tensor, alpha: float):
    """One elastic round (lua/AllReduceEA.lua:34-45): returns (new_xs, new_center)."""
    alpha_t = torch.tensor(alpha, dtype=center.dtype)
    total = torch.zeros_like(center)
    new_xs = []
    for x in xs:
        d = (x - center) * alpha_t
        new_xs.append(x - d)
        total = total + d
    return new_xs, center + total


def winner(steps: Sequence[int]) -> int:
    """Node whose parameters win synchronizeParameters: most steps, ties -> highest index
    (ascending sort, last element: lua/AllReduceSGD.lua:41-43); everyone at 0 -> node 1 (scatter, :52)."""
    best = max(steps)
    if best == 0:
        return 0
    return max(r for r, s in enumerate(steps) if s == best)


class SgdNodeSim:
    """N nodes running the reference's AllReduceSGD unit test body (test/test_AllReduceSGD.lua:8-21)
    with per-node step counts; API-parity semantics: every node applies its own update, a finished
    node stops updating, the winner's parameters are broadcast at the end of the epoch."""

    def __init__(self, params: List[torch.Tensor]):
        self.params = [p.clone() for p in params]
        self.n = len(params)

    def synchronize(self, steps: Sequence[int]) -> int:
        w = winner(steps)
        self.params = [self.params[w].clone() for _ in range(self.n)]
        return w

    def epoch(self, steps_per_node: Sequence[int], grad_fn) -> None:
        """grad_fn(node, step_index, steps_of_node) -> gradient tensor."""
        for s in range(max(steps_per_node)):
            active = [1 if s < k else 0 for k in steps_per_node]
            grads = [grad_fn(r, s, steps_per_node[r]) if active[r] else torch.zeros_like(self.params[r]) for r in range(self.n)]
            g, _ = allreduce_sum(grads, active, normalize=True)
            for r in range(self.n):
                if active[r]:
                    self.params[r] = self.params[r] + g   # the unit test does params:add(grads)
        self.synchronize(steps_per_node)


class EaNodeSim:
    """N nodes of AllReduceEA with uneven steps (lua/AllReduceEA.lua) under this library's drain
    spec: a node that finished keeps taking part in the others' elastic rounds; the terminating
    round moves nothing."""

    def __init__(self, xs: List[torch.Tensor], tau: int, alpha: float):
        self.xs = [x.clone() for x in xs]
        self.n = len(xs)
        self.tau, self.alpha = tau, alpha
        self.center = None
        self.step = [0] * self.n

    def synchronize_parameters(self) -> None:
        self.xs = [self.xs[0].clone() for _ in range(self.n)]
        self.center = self.xs[0].clone()
        self.step = [0] * self.n

    def epoch(self, steps_per_node: Sequence[int], move_fn) -> None:
        """move_fn(node, step_index) -> perturbation added to x before averageParameters."""
        rounds_needed = [k // self.tau for k in [s + st for s, st in zip(steps_per_node, self.step)]]
        done_rounds = 0
        local = list(self.step)
        remaining = list(steps_per_node)
        idx = [0] * self.n
        total_rounds = max((local[r] + steps_per_node[r]) // self.tau - local[r] // self.tau for r in range(self.n))
        for _ in range(total_rounds):
            # every node that still has steps advances to its next multiple of tau (or runs out)
            for r in range(self.n):
                while remaining[r] > 0:
                    self.xs[r] = self.xs[r] + move_fn(r, idx[r])
                    idx[r] += 1
                    remaining[r] -= 1
                    local[r] += 1
                    if local[r] % self.tau == 0:
                        break
            self.xs, self.center = elastic_average(self.xs, self.center, self.alpha)
        # leftover local steps that never reached a multiple of tau
        for r in range(self.n):
            while remaining[r] > 0:
                self.xs[r] = self.xs[r] + move_fn(r, idx[r])
                idx[r] += 1
                remaining[r] -= 1
        self.step = [0] * self.n


def async_client_sync(x: torch.Tensor, center: torch.Tensor, alpha: float):
    """One client sync (lua/AsyncEA.lua:109-119 + :212-216): returns (new_x, new_center)."""
    d = (x - center) * torch.tensor(alpha, dtype=x.dtype)
    return x - d, center + d
