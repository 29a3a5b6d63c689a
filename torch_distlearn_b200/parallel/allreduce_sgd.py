"""AllReduceSGD — synchronous data-parallel SGD (reference: lua/AllReduceSGD.lua).

API parity (camelCase names are the reference's; snake_case aliases are provided):

* ``sumGradients(grads)``               lua/AllReduceSGD.lua:10-15
* ``sumAndNormalizeGradients(grads)``   lua/AllReduceSGD.lua:18-30   (returns n contributors)
* ``synchronizeParameters(params)``     lua/AllReduceSGD.lua:33-54

plus the Blackwell-native fused path

* ``step(flat, lr, ...)`` — ONE kernel launch that all-reduces the gradients of the active ranks,
  divides by their number, applies (momentum / weight-decay) SGD and distributes the new
  parameters: the reference's ``sumAndNormalizeGradients`` + the user's ``p:add(-lr, g)`` loop
  (examples/cifar10.lua:184-191).

Uneven step counts (a rank whose data partition has fewer batches) are handled like the reference:
a rank that finished its epoch enters ``synchronizeParameters`` and *drains* — it keeps answering
the others' collectives with ``active = 0`` (contributing zeros, lua/AllReduceSGD.lua:37) until a
round in which nobody is active; then the rank with the most steps (ties: highest index, the
last element of the ascending sort at :41-43) broadcasts its parameters.  Differences from the
reference, on purpose: the decision to drain is collective (the reference branches on a
per-node counter, :35, and can deadlock when one node has an empty partition), and step counts
travel in the barrier payload instead of a separate all-reduce (:39).
"""
from __future__ import annotations

from typing import Any, Optional

from .comm import Communicator
from .flat import FlatParams, TableStager


class AllReduceSGD:
    def __init__(self, comm: Communicator, flat: Optional[FlatParams] = None, channel: int = 0):
        self.comm = comm
        self.flat = flat
        self.channel = channel
        self.steps = 0                      # stepsPerNode[self] (lua/AllReduceSGD.lua:7)
        self._stager = TableStager(comm)
        self._last = None                   # geometry + kind of the last collective step (for draining)
        self.last_n = 0
        if flat is not None and flat.grads is not None:
            self._last = ("allreduce", flat.grads, flat.padded, None)

    # ------------------------------------------------------------------------------------------------
    def _reduce(self, grads: Any, normalize: bool) -> int:
        buf, n, tensors = self._stager.resolve(grads, "grads")
        if tensors is not None:
            self._stager.pack(buf, tensors)
        res = self.comm.allreduce(buf, n, active=1, payload=min(self.steps + 1, 2**30 - 1), normalize=normalize,
                                  channel=self.channel, read=self._need_n)
        if tensors is not None:
            self._stager.unpack(buf, tensors)
        self._last = ("allreduce", buf, n, None)
        self.steps += 1
        if res is not None:
            self.last_n = res.n_active
        elif hasattr(self.comm, "lazy_count"):
            # the kernel leaves n in pinned host memory; the value is fetched (event wait + plain load) when used
            self.last_n = self.comm.lazy_count(self.channel, self.comm.last_ticket)
        else:
            self.last_n = self.comm.world
        return self.last_n

    # On the GPU backends the count comes back as a LazyCount: exact, but read only when it is used, so a caller
    # that ignores it (every caller in the reference's examples) never synchronises.  Set ``want_contributors = True``
    # to get a plain int from every call (one host synchronisation per call).
    want_contributors = False

    @property
    def _need_n(self) -> bool:
        return self.want_contributors or self.comm.backend == "host"

    def sumGradients(self, grads: Any) -> int:
        """grads <- sum over contributing nodes (lua/AllReduceSGD.lua:10-15)."""
        return self._reduce(grads, normalize=False)

    def sumAndNormalizeGradients(self, grads: Any) -> int:
        """grads <- sum over contributing nodes / #contributors (lua/AllReduceSGD.lua:18-30)."""
        return self._reduce(grads, normalize=True)

    # ------------------------------------------------------------------------------------------------
    def step(self, flat: Optional[FlatParams] = None, lr: float = 0.01, momentum: float = 0.0,
             weight_decay: float = 0.0, nesterov: bool = False, zero_grads: bool = True, variant: str = "auto",
             stream=None, nblocks: int = 0):
        """Fused all-reduce + normalise + SGD update + parameter all-gather (one launch)."""
        flat = flat or self.flat
        if flat is None or flat.grads is None:
            raise ValueError("step() needs FlatParams with gradients")
        if momentum != 0.0 and flat.momentum is None:
            raise ValueError("momentum needs FlatParams(with_momentum=True)")
        kw = dict(momentum=momentum, weight_decay=weight_decay, nesterov=nesterov, zero_grads=zero_grads, variant=variant)
        extra = {}
        if self.comm.backend != "host":
            extra = dict(stream=stream, nblocks=nblocks)
        self.comm.fused_sgd(flat.grads, flat.params, flat.momentum if momentum != 0.0 else None, flat.padded, lr,
                            active=1, payload=min(self.steps + 1, 2**30 - 1), channel=self.channel, read=False, **kw, **extra)
        self._last = ("fused", flat, flat.padded, dict(lr=lr, **kw))
        self.steps += 1

    # ------------------------------------------------------------------------------------------------
    def _drain(self) -> list:
        """Serve the stragglers' collectives with zero contributions until nobody is active;
        returns every rank's step count (lua/AllReduceSGD.lua:37-39)."""
        payload = min(self.steps, 2**30 - 1)
        while True:
            if self._last is None:
                res = self.comm.barrier(active=0, payload=payload, channel=self.channel, read=True)
            elif self._last[0] == "bucketed":
                res = self._last[1].drain_round()
            elif self._last[0] == "fused":
                _, flat, n, kw = self._last
                self.comm.fused_sgd(flat.grads, flat.params, flat.momentum if kw["momentum"] != 0.0 else None, n,
                                    kw["lr"], momentum=kw["momentum"], weight_decay=kw["weight_decay"],
                                    nesterov=kw["nesterov"], zero_grads=kw["zero_grads"], variant=kw["variant"],
                                    active=0, payload=payload, channel=self.channel, read=False)
                res = self.comm.read_result(self.channel)
            else:
                _, buf, n, _ = self._last
                self.comm.allreduce(buf, n, active=0, payload=payload, normalize=True, channel=self.channel, read=False)
                res = self.comm.read_result(self.channel)
            if res.n_active == 0:
                return res.payloads

    def synchronizeParameters(self, params: Any = None) -> int:
        """Every node ends with bit-identical copies of the parameters of the node that took the most
        steps this epoch (lua/AllReduceSGD.lua:33-54); returns that node's 0-based rank."""
        params = params if params is not None else self.flat
        buf, n, tensors = self._stager.resolve(params, "params")
        if self._last is None:
            # learn the gradient geometry from the parameters (same shapes) so that a rank with an
            # empty partition can still serve its peers' all-reduces
            if tensors is not None:
                gbuf, gn, _ = self._stager.resolve(params, "grads")
                self._last = ("allreduce", gbuf, gn, None)
            elif isinstance(params, FlatParams) and params.grads is not None:
                self._last = ("allreduce", params.grads, params.padded, None)
        steps = self._drain()
        # most steps wins; ties -> highest index (stepsPerNode:sort() ascending, take the last)
        best = max(steps)
        root = max(r for r, s in enumerate(steps) if s == best) if best > 0 else 0
        if tensors is not None:
            self._stager.pack(buf, tensors)
        self.comm.broadcast(buf, n, root=root, channel=self.channel, read=False)
        if tensors is not None:
            self._stager.unpack(buf, tensors)
        self.steps = 0
        return root

    # snake_case aliases
    sum_gradients = sumGradients
    sum_and_normalize_gradients = sumAndNormalizeGradients
    synchronize_parameters = synchronizeParameters
