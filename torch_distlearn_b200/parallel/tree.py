"""``Tree`` / ``LocalhostTree`` facade — the reference's transport handle
(``ipc.Tree(nodeIndex, numNodes, base, server, client, host, port)``,
test/test_AllReduceSGD.lua:7; ``ipc.LocalhostTree(nodeIndex, numNodes)``, examples/mnist.lua:16).

Scripts written against the reference call ``tree.allReduce(value, reduceFn[, contribFn])``,
``tree.scatter(value)``, ``tree.walkTable`` and read ``tree.nodeIndex`` / ``tree.numNodes``
(examples/cifar10.lua:203: the confusion-matrix reduce).  This class offers the same surface on
top of a :class:`Communicator`.  There is no tree any more: behind an NVSwitch every GPU reaches
every peer at full bandwidth, so the collectives are direct one-/two-shot/NVLS kernels, not
log-depth reduce-up/broadcast-down (lua/AllReduceEA.md:26-30).

Only summation is supported as ``reduceFn`` (every call site in the reference passes
``function(a, b) return a:add(b) end``); the argument is accepted and ignored.
"""
from __future__ import annotations

import os
from typing import Any, Callable, Optional, Tuple

import torch

from .comm import Communicator, HostCommunicator, P2PCommunicator
from .flat import TableStager, walk_table


class Tree:
    def __init__(self, comm: Communicator, channel: int = 1):
        self.comm = comm
        self.nodeIndex = comm.node_index       # 1-based, node 1 is the root
        self.numNodes = comm.num_nodes
        self.channel = channel
        self._stager = TableStager(comm)

    node_index = property(lambda self: self.nodeIndex)
    num_nodes = property(lambda self: self.numNodes)

    @staticmethod
    def walkTable(table: Any, fn: Callable[[torch.Tensor], Any]) -> None:
        walk_table(table, fn)

    def allReduce(self, value: Any, reduceFn: Optional[Callable] = None, contribFn: Optional[Callable] = None,
                  like: Any = None) -> Tuple[Any, int]:
        """In-place sum of ``value`` over the nodes; returns ``(value, n)`` with n = number of nodes
        that contributed actively.  With ``contribFn`` the caller is *finished*: it serves rounds,
        contributing ``contribFn(tensor, i)`` for the i-th tensor (1-based), until a round in which
        nobody is active (lua/AllReduceSGD.lua:37, lua/AllReduceEA.lua:58-68).  ``value=None`` with a
        contribFn needs ``like=`` (a table of the right geometry) — the reference gets the geometry
        from the wire."""
        if contribFn is None:
            buf, n, tensors = self._stager.resolve(value, "value")
            if tensors is not None:
                self._stager.pack(buf, tensors)
            res = self.comm.allreduce(buf, n, active=1, normalize=False, channel=self.channel, read=True)
            if tensors is not None:
                self._stager.unpack(buf, tensors)
            return value, res.n_active
        table = value if value is not None else like
        if table is None:
            raise ValueError("tree.allReduce(None, fn, contribFn) needs like=<table with the peers' geometry>")
        buf, n, tensors = self._stager.resolve(table, "value")
        n_last = 0
        while True:
            if tensors is not None:
                for i, t in enumerate(tensors, 1):
                    out = contribFn(t, i)
                    if out is not None and out is not t:
                        t.copy_(out)
                self._stager.pack(buf, tensors)
            # active=2: a finished node's contribution is in the sum but it is not counted in n
            res = self.comm.allreduce(buf, n, active=2, normalize=False, channel=self.channel, read=True)
            if tensors is not None:
                self._stager.unpack(buf, tensors)
            n_last = res.n_active
            if n_last == 0:
                break
        return value, n_last

    def scatter(self, value: Any) -> Any:
        """Node 1's value overwrites everyone's (lua/AllReduceSGD.lua:50-52)."""
        buf, n, tensors = self._stager.resolve(value, "scatter")
        if buf.dtype != torch.float32:
            raise TypeError("tree.scatter supports float32 tables")
        if tensors is not None:
            self._stager.pack(buf, tensors)
        self.comm.broadcast(buf, n, root=0, channel=self.channel, read=False)
        if tensors is not None:
            self._stager.unpack(buf, tensors)
        return value

    all_reduce = allReduce
    walk_table = walkTable


def LocalhostTree(nodeIndex: int, numNodes: int, port: int = 8080, backend: str = "auto", heap_mb: float = 64) -> Tree:
    """``ipc.LocalhostTree(nodeIndex, numNodes)`` (examples/mnist.lua:16): all nodes on this machine.
    CPU jobs rendezvous on a shared-memory segment named after ``port``; GPU jobs expect
    ``torch.distributed`` to be initialised (see ``launch.init_from_env``)."""
    if backend == "auto":
        backend = "p2p" if torch.cuda.is_available() else "host"
    if backend == "host":
        name = f"distlearn-{os.environ.get('DISTLEARN_JOB', os.getuid())}-{port}"
        comm = HostCommunicator(nodeIndex - 1, numNodes, name=name, heap_mb=heap_mb)
    else:
        comm = P2PCommunicator(rank=nodeIndex - 1, world=numNodes, heap_mb=heap_mb)
    return Tree(comm)
