"""AllReduceEA — synchronous Elastic Averaging SGD (reference: lua/AllReduceEA.lua, lua/AllReduceEA.md).

Every ``tau`` local steps each node moves towards a shared center point and the center moves
towards the nodes:

    delta_i = alpha * (x_i - c) ;  x_i -= delta_i ;  c += sum_i delta_i        (AllReduceEA.md:12-24)

The reference runs that as three elementwise passes around a tree all-reduce of ``delta``; here it
is ONE kernel (``csrc/fused_ea.cu``): slice owners read peers' ``x`` directly over NVLink, the new
center is pushed to every replica (double buffered), ``x`` is moved locally.  ``delta`` is never
materialised, and center replicas stay bit-identical instead of drifting (cf. the comment at
lua/AllReduceEA.lua:76).

API parity: ``averageParameters`` (:25-47), ``synchronizeCenter`` (:77-84),
``synchronizeParameters`` (:87-100).  Uneven step counts drain as in the reference (:50-72): a
finished node keeps taking part in the others' elastic rounds — contributing its own elastic move,
not zeros — until a round in which nobody is active; that terminating round moves nothing.
"""
from __future__ import annotations

from typing import Any, Optional

import torch

from .comm import Communicator, SymBuffer
from .flat import FlatParams, TableStager


class AllReduceEA:
    def __init__(self, comm: Communicator, tau: int, alpha: float, flat: Optional[FlatParams] = None, channel: int = 0):
        if tau < 1:
            raise ValueError("tau must be >= 1")
        self.comm, self.tau, self.alpha, self.channel = comm, int(tau), float(alpha), channel
        self.flat = flat
        self.step = 0                       # lua/AllReduceEA.lua:5
        self._stager = TableStager(comm)
        self._x: Optional[SymBuffer] = None
        self._n = 0
        self._center = [None, None]         # double-buffered center replicas
        self._cur = 0
        if flat is not None:
            self._one_time_init(flat)

    # ------------------------------------------------------------------------------------------------
    def _one_time_init(self, params: Any):
        """Clone the parameters into the center point (lua/AllReduceEA.lua:11-22)."""
        buf, n, tensors = self._stager.resolve(params, "params")
        if self._x is None:
            self._x, self._n = buf, n
            self._center = [self.comm.alloc(n, torch.float32, "ea.center0"), self.comm.alloc(n, torch.float32, "ea.center1")]
            if tensors is not None:
                self._stager.pack(buf, tensors)
            self._center[0].tensor[:n].copy_(buf.tensor[:n])
            self._cur = 0
        elif buf.offset != self._x.offset or n != self._n:
            raise ValueError("AllReduceEA was initialised with a different parameter table")
        return buf, n, tensors

    @property
    def center(self) -> torch.Tensor:
        """The local replica of the center point (flat view)."""
        return self._center[self._cur].tensor

    def _launch(self, active: int, read: bool):
        res = self.comm.fused_ea(self._x, self._center[self._cur], self._center[1 - self._cur], self._n, self.alpha,
                                 active=active, payload=min(self.step, 2**30 - 1), channel=self.channel, read=read)
        self._cur = 1 - self._cur           # every rank flips in lockstep (the kernel keeps c_new coherent)
        return res

    # ------------------------------------------------------------------------------------------------
    def averageParameters(self, params: Any = None) -> bool:
        """Call once per local step; every ``tau``-th call runs the elastic average
        (lua/AllReduceEA.lua:25-47).  Returns True when it did."""
        params = params if params is not None else self.flat
        buf, n, tensors = self._one_time_init(params)
        self.step += 1
        if self.step % self.tau != 0:
            return False
        if tensors is not None:
            self._stager.pack(buf, tensors)
        self._launch(active=1, read=False)
        if tensors is not None:
            self._stager.unpack(buf, tensors)
        return True

    def _handle_uneven_steps(self, tensors) -> None:
        """Drain (lua/AllReduceEA.lua:50-72).  The decision is collective: every rank runs at least
        the terminating round, so a rank with zero steps cannot diverge from its peers."""
        if tensors is not None:
            self._stager.pack(self._x, tensors)
        while True:
            self._launch(active=0, read=False)
            res = self.comm.read_result(self.channel)
            if res.n_active == 0:
                break
        if tensors is not None:
            self._stager.unpack(self._x, tensors)
        self.step = 0

    def synchronizeCenter(self, params: Any = None) -> None:
        """Same center on every node (lua/AllReduceEA.lua:77-84): drain, then broadcast node 1's."""
        params = params if params is not None else self.flat
        _, _, tensors = self._one_time_init(params)
        self._handle_uneven_steps(tensors)
        self.comm.broadcast(self._center[self._cur], self._n, root=0, channel=self.channel, read=False)

    def synchronizeParameters(self, params: Any = None) -> None:
        """Same parameters on every node and center <- parameters (lua/AllReduceEA.lua:87-100)."""
        params = params if params is not None else self.flat
        buf, n, tensors = self._one_time_init(params)
        self._handle_uneven_steps(tensors)
        if tensors is not None:
            self._stager.pack(buf, tensors)
        self.comm.broadcast(buf, n, root=0, also=self._center[self._cur], channel=self.channel, read=False)
        if tensors is not None:
            self._stager.unpack(buf, tensors)

    average_parameters = averageParameters
    synchronize_center = synchronizeCenter
    synchronize_parameters = synchronizeParameters
