"""Flat (getParameters()-style) parameter / gradient vectors in the symmetric heap.

The reference keeps ``params``/``grads`` as *tables of tensors* whose storage must stay put so
the transport can register it once (``stableGradients = true  -- ... so we can use CUDA IPC``,
examples/mnist.lua:91-94).  Here the idea is taken to its end: one flat fp32 buffer for all
parameters and one for all gradients, allocated in the symmetric heap, with every
``nn.Parameter`` (and its ``.grad``) re-pointed to a view of it.  The collectives then run
in place on addresses every peer already has mapped — no packing, no staging copies.

Also here: the table walker (``ipc.utils.walkTable``, lua/AsyncEA.lua:3,23) and the staging
path that lets the algorithm classes accept arbitrary tables of tensors for API parity.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, Iterable, List, Optional, Sequence, Tuple, Union

import torch

from .comm import Communicator, SymBuffer


def walk_table(table: Any, fn: Optional[Callable[[torch.Tensor], Any]] = None) -> List[torch.Tensor]:
    """Visit every tensor of a nested list/tuple/dict in deterministic order (list order, dict
    insertion order) — the counterpart of ``tree.walkTable`` (lua/AllReduceSGD.lua:24)."""
    out: List[torch.Tensor] = []

    def rec(x):
        if isinstance(x, torch.Tensor):
            out.append(x)
            if fn is not None:
                fn(x)
        elif isinstance(x, FlatParams):
            rec(x.params.tensor)
        elif isinstance(x, SymBuffer):
            rec(x.tensor)
        elif isinstance(x, dict):
            for k in x:
                rec(x[k])
        elif isinstance(x, (list, tuple)):
            for v in x:
                rec(v)
        elif isinstance(x, torch.nn.Module):
            for p in x.parameters():
                rec(p.data)
        elif x is None:
            return
        else:
            raise TypeError(f"walk_table: unsupported leaf {type(x)!r}")

    rec(table)
    return out


class FlatParams:
    """Flattens a module's (or a list of tensors') parameters and gradients into symmetric
    buffers.  ``params`` / ``grads`` (/ ``momentum``) are :class:`SymBuffer`; ``numel`` is the
    true element count and ``padded`` the allocated (32-element aligned) one.

    All ranks must construct FlatParams for the same architecture in the same order.

    ``order`` is the layout of the flat vector: ``"module"`` (``named_parameters()`` order, what ``getParameters()``
    gives the reference), ``"execution"`` (the order in which the forward pass uses the parameters, taken from the
    module's ``execution_order()`` if it has one), or an explicit list of parameter names.  Execution order makes the
    tail of the vector the first thing backward finishes, so :class:`GradBuckets` can send buckets off layer by layer;
    in module order a bucket waits for the LAST of its tensors (e.g. every BatchNorm of a model that registers all its
    convolutions first).  Checkpoints record the layout and are remapped by name on load."""

    def __init__(self, comm: Communicator, source: Union[torch.nn.Module, Iterable[torch.Tensor]],
                 with_grads: bool = True, with_momentum: bool = False, name: str = "model",
                 order: Union[str, Sequence[str]] = "module"):
        self.comm = comm
        if isinstance(source, torch.nn.Module):
            named = list(source.named_parameters())
            if order == "execution":
                order = source.execution_order() if hasattr(source, "execution_order") else "module"
            if order != "module":
                if isinstance(order, str):
                    raise ValueError(f"FlatParams: unknown order {order!r}")
                by_name = dict(named)
                if sorted(order) != sorted(by_name):
                    raise ValueError("FlatParams: order must list every parameter name exactly once")
                named = [(n, by_name[n]) for n in order]
            self.names = [n for n, _ in named]
            self._params: List[torch.Tensor] = [p for _, p in named]
        else:
            if order != "module":
                raise ValueError("FlatParams: order needs a module (a list of tensors is taken as given)")
            self._params = list(source)
            self.names = [f"p{i}" for i in range(len(self._params))]
        for p in self._params:
            if p.dtype != torch.float32:
                raise TypeError("FlatParams: parameters must be float32 (the reference casts everything to float, "
                                "examples/cifar10.lua:136)")
        self.shapes = [tuple(p.shape) for p in self._params]
        self.sizes = [p.numel() for p in self._params]
        self.offsets: List[int] = []
        cur = 0
        for s in self.sizes:
            self.offsets.append(cur)
            cur += (s + 7) // 8 * 8  # every tensor starts on a 32-byte boundary: any range of whole tensors (a gradient
            # bucket) can then go through the 256-bit paths of the collectives
        self.numel = cur
        self.padded = comm.padded(cur)
        self.params = comm.alloc(self.padded, torch.float32, name + ".params")
        self.grads = comm.alloc(self.padded, torch.float32, name + ".grads") if with_grads else None
        self.momentum = comm.alloc(self.padded, torch.float32, name + ".momentum") if with_momentum else None
        self.param_views: List[torch.Tensor] = []
        self.grad_views: List[torch.Tensor] = []
        with torch.no_grad():
            for p, off, size, shape in zip(self._params, self.offsets, self.sizes, self.shapes):
                pv = self.params.tensor[off:off + size].view(shape)
                pv.copy_(p.detach().to(pv.device))
                self.param_views.append(pv)
                if isinstance(p, torch.nn.Parameter) or p.requires_grad:
                    p.data = pv
                if with_grads:
                    gv = self.grads.tensor[off:off + size].view(shape)
                    self.grad_views.append(gv)
                    if isinstance(p, torch.nn.Parameter) or p.requires_grad:
                        p.grad = gv
        self.device = self.params.tensor.device

    # gradients are views: never drop them, zero in place (the fused step does it in-kernel)
    def zero_grad(self) -> None:
        if self.grads is not None:
            self.grads.tensor.zero_()

    def attach_grads(self) -> None:
        """Re-point ``p.grad`` at the flat views (after something replaced or dropped them)."""
        for p, gv in zip(self._params, self.grad_views):
            if p.grad is None or p.grad.data_ptr() != gv.data_ptr():
                if p.grad is not None:
                    gv.copy_(p.grad)
                p.grad = gv

    def manifest(self) -> List[Dict[str, Any]]:
        return [{"name": n, "shape": list(s), "offset": o, "numel": z}
                for n, s, o, z in zip(self.names, self.shapes, self.offsets, self.sizes)]

    def __repr__(self) -> str:
        return f"FlatParams({len(self.sizes)} tensors, numel={self.numel}, padded={self.padded}, device={self.device})"


class TableStager:
    """Lets the API-parity entry points accept any table of float tensors: if the argument already
    IS a symmetric buffer it is used in place, otherwise it is packed into a cached staging
    buffer before the collective and unpacked after."""

    def __init__(self, comm: Communicator):
        self.comm = comm
        self._cache: Dict[Tuple[str, int, torch.dtype], SymBuffer] = {}

    def resolve(self, table: Any, role: str) -> Tuple[SymBuffer, int, Optional[List[torch.Tensor]]]:
        """Returns (buffer, n, tensors_to_unpack_or_None)."""
        if isinstance(table, FlatParams):
            buf = table.grads if role == "grads" and table.grads is not None else table.params
            return buf, buf.numel, None
        if isinstance(table, SymBuffer):
            return table, table.numel, None
        tensors = walk_table(table)
        if not tensors:
            raise ValueError("empty table")
        dtype = tensors[0].dtype
        total = sum((t.numel() + 3) // 4 * 4 for t in tensors)
        es = tensors[0].element_size()
        padded = (total * es + 15) // 16 * 16 // es
        padded = max(padded, 16 // es)
        if dtype == torch.float32:
            padded = self.comm.padded(total)
        key = (role, padded, dtype)
        buf = self._cache.get(key)
        if buf is None:
            # Heap offsets must agree across ranks, and ranks do not meet the roles in the same order: a rank whose
            # partition was empty reaches "params" (synchronizeParameters) before it ever staged a gradient, its peers
            # the other way round.  So the gradient and parameter staging buffers of one geometry are always carved
            # out together, in one fixed order, whichever role is asked for first.
            group = self._PAIRED if role in self._PAIRED else (role,)
            for r in group:
                k = (r, padded, dtype)
                if k not in self._cache:
                    self._cache[k] = self.comm.alloc(padded, dtype, f"stage.{r}")
            buf = self._cache[key]
        return buf, padded, tensors

    _PAIRED = ("grads", "params")

    @staticmethod
    def pack(buf: SymBuffer, tensors: Sequence[torch.Tensor]) -> None:
        off = 0
        flat = buf.tensor
        for t in tensors:
            n = t.numel()
            flat[off:off + n].copy_(t.reshape(-1), non_blocking=True)
            off += (n + 3) // 4 * 4

    @staticmethod
    def unpack(buf: SymBuffer, tensors: Sequence[torch.Tensor]) -> None:
        off = 0
        flat = buf.tensor
        for t in tensors:
            n = t.numel()
            t.copy_(flat[off:off + n].view(t.shape), non_blocking=True)
            off += (n + 3) // 4 * 4
