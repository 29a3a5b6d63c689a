"""TrainStep — the whole synchronous training step as ONE CUDA-graph replay.

    host batch (pinned) --H2D--> static input --> forward --> loss --> backward (grads land in the
    flat symmetric gradient buffer) --> fused all-reduce + 1/n + SGD + param all-gather kernel

The reference's step is ``df(params, x, y)`` -> ``sumAndNormalizeGradients(grads)`` -> a Lua loop of
``p:add(-lr, g)`` (examples/cifar10.lua:181-191): dozens of small launches plus a socket tree.
Here the forward/backward kernels (cuDNN/cuBLAS through PyTorch) and the fused collective kernel are
captured once and replayed, so a step costs one graph launch on the host; the cross-GPU barriers
inside the collective are device-side flags with device-side epochs, which is what makes the
kernel capturable.  Every rank must replay the same number of steps between synchronisation points
(use ``AllReduceSGD.synchronizeParameters`` / the eager path for uneven partitions).
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch
import torch.nn.functional as F

from .allreduce_sgd import AllReduceSGD
from .flat import FlatParams


class TrainStep:
    def __init__(self, model: torch.nn.Module, flat: FlatParams, sgd: AllReduceSGD, batch_shape: Tuple[int, ...],
                 lr: float, momentum: float = 0.0, weight_decay: float = 0.0, nesterov: bool = False,
                 loss_fn: Optional[Callable] = None, use_graph: bool = True, variant: str = "auto",
                 target_dtype: torch.dtype = torch.int64, warmup: int = 3, nblocks: int = 0, bucket_mb: float = 0.0,
                 bucket_blocks: int = 0, restore_warmup: bool = True):
        self.model, self.flat, self.sgd = model, flat, sgd
        self.lr, self.momentum, self.weight_decay, self.nesterov = lr, momentum, weight_decay, nesterov
        self._fused_loss = loss_fn is None and hasattr(model, "forward_with_loss")
        self.loss_fn = loss_fn or F.nll_loss          # the models end in log_softmax (ClassNLLCriterion)
        self.variant, self.nblocks = variant, nblocks
        dev = flat.device
        self.device = dev
        self.x = torch.zeros(batch_shape, dtype=torch.float32, device=dev)
        self.y = torch.zeros(batch_shape[0], dtype=target_dtype, device=dev)
        self.loss = torch.zeros((), dtype=torch.float32, device=dev)
        self.logits: Optional[torch.Tensor] = None
        self.loss_host = torch.zeros((), dtype=torch.float32).pin_memory() if dev.type == "cuda" else torch.zeros(())
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.launches_per_step = 0
        self.kernels_per_step = 0
        self._replayed_kernels = 0
        self.use_graph = bool(use_graph and dev.type == "cuda")
        self._warmup = warmup
        self._restore_warmup = restore_warmup
        self._captured = False
        self._copy_stream = None
        self._staged = False
        # bucket_mb > 0: launch the fused kernel per gradient bucket from autograd hooks, overlapped
        # with the rest of backward on a side stream (parallel/bucketed.py)
        self.buckets = None
        if bucket_mb > 0:
            from .bucketed import GradBuckets
            self.buckets = GradBuckets(flat, sgd, bucket_mb=bucket_mb, nblocks=bucket_blocks, variant=variant)
            self.buckets.set_hyper(lr, momentum, weight_decay, nesterov)

    # ------------------------------------------------------------------------------------------------
    def _forward_backward(self) -> None:
        if self._fused_loss:
            out, loss = self.model.forward_with_loss(self.x, self.y)     # classifier head + loss in the model's own kernels
        else:
            out = self.model(self.x)
            loss = self.loss_fn(out, self.y)
        loss.backward()
        self.logits = out.detach()
        self.loss.copy_(loss.detach())

    def _body(self) -> None:
        self._forward_backward()
        if self.buckets is not None:
            self.buckets.finish()           # the per-bucket kernels were launched from the backward hooks
        else:
            self.sgd.step(self.flat, lr=self.lr, momentum=self.momentum, weight_decay=self.weight_decay,
                          nesterov=self.nesterov, zero_grads=True, variant=self.variant, nblocks=self.nblocks)

    def _snapshot(self):
        flat = self.flat
        return (flat.params.tensor.clone(), flat.momentum.tensor.clone() if flat.momentum is not None else None,
                [b.clone() for b in self.model.buffers()], self.sgd.steps)

    def _restore(self, saved) -> None:
        params, momentum, buffers, steps = saved
        with torch.no_grad():
            self.flat.params.tensor.copy_(params)       # identical on every rank before and after: nothing to re-sync
            if momentum is not None:
                self.flat.momentum.tensor.copy_(momentum)
            for b, old in zip(self.model.buffers(), buffers):
                b.copy_(old)
        self.sgd.steps = steps

    def _native_count(self) -> int:
        if self.device.type != "cuda":
            return self.sgd.comm.launches
        from .. import _lib
        return int(_lib.cuda().dl_kernel_launches())

    def _eager(self) -> None:
        before, nbefore = self.sgd.comm.launches, self._native_count()
        self._body()
        self.launches_per_step = self.sgd.comm.launches - before
        self.kernels_per_step = self._native_count() - nbefore      # collectives + tensor-core convolutions + helpers

    @property
    def kernel_launches(self) -> int:
        """Kernels of this library launched so far (host-counted launches + kernels replayed inside CUDA graphs)."""
        return self._native_count() + self._replayed_kernels

    def capture(self) -> None:
        """Warm up eagerly on the current contents of the static buffers (loads modules, lets cuDNN
        pick algorithms), then capture one step when graphs are on.  The warm-up steps are full steps — every rank
        launches the collectives — but they leave no trace: parameters, momentum, the module's buffers (BatchNorm
        running statistics) and the step counter are put back afterwards, so training starts from the state the caller
        prepared.  ``step(x, y)`` loads the batch first, so the warm-up sees real data; the buffers start as zeros."""
        if self._captured:
            return
        self._captured = True
        self.flat.attach_grads()
        if self.device.type != "cuda":
            return
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            saved = self._snapshot() if self._restore_warmup else None
            for _ in range(max(self._warmup, 1)):
                self._eager()
            if saved is not None:
                self._restore(saved)
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        if not self.use_graph:
            return
        self._recapture()              # the capture itself does not execute a step

    # ------------------------------------------------------------------------------------------------
    def load(self, x: torch.Tensor, y: torch.Tensor) -> None:
        """Copy a batch into the static input buffers (async H2D when the source is pinned)."""
        self.x.copy_(x, non_blocking=True)
        self.y.copy_(y, non_blocking=True)

    def prefetch(self, x: torch.Tensor, y: torch.Tensor) -> None:
        """Start copying the NEXT batch (pinned host memory) to the device on a copy stream while the current step
        is still running; the following ``run()`` picks it up.  The input pipeline of a real trainer: the H2D copy
        disappears behind the step instead of sitting in front of it."""
        if self.device.type != "cuda":
            self.load(x, y)
            return
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(self.device)
            self._x_next, self._y_next = torch.empty_like(self.x), torch.empty_like(self.y)
            self._staged_ev, self._consumed_ev = torch.cuda.Event(), torch.cuda.Event()
            self._consumed_ev.record(torch.cuda.current_stream(self.device))
        self._copy_stream.wait_event(self._consumed_ev)          # the previous staged batch has been moved out
        with torch.cuda.stream(self._copy_stream):
            self._x_next.copy_(x, non_blocking=True)
            self._y_next.copy_(y, non_blocking=True)
            self._staged_ev.record(self._copy_stream)
        self._staged = True

    def run(self) -> torch.Tensor:
        """One training step on whatever is in the static buffers (or on the prefetched batch); returns the (device) loss."""
        if not self._captured:
            self.capture()
        if self._staged:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(self._staged_ev)
            self.x.copy_(self._x_next, non_blocking=True)        # device-to-device, a few microseconds
            self.y.copy_(self._y_next, non_blocking=True)
            self._consumed_ev.record(cur)
            self._staged = False
        if self.graph is not None:
            self.graph.replay()
            self.sgd.steps += 1
            self.sgd.comm.launches += self.launches_per_step
            self._replayed_kernels += self.kernels_per_step
        else:
            self._eager()
        return self.loss

    def step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """The user-facing call: load a (host or device) batch, run the step, return the device loss."""
        self.load(x, y)
        return self.run()

    def loss_value(self) -> float:
        """Device->host read of the last step's loss (synchronises).  Also the place where a collective that gave up
        waiting for a peer surfaces: the sticky error word in pinned host memory is checked here (a plain load)."""
        if self.device.type == "cuda":
            self.loss_host.copy_(self.loss, non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
            poll = getattr(self.sgd.comm, "poll_error", None)
            if poll is not None:
                poll()
            return float(self.loss_host)
        return float(self.loss)

    def set_lr(self, lr: float, momentum: Optional[float] = None, weight_decay: Optional[float] = None) -> None:
        """Change the hyper-parameters of later steps (a learning-rate schedule).  They are kernel arguments, so a
        captured graph has them baked in: the graph is dropped and captured again on the next ``run()`` (a few
        milliseconds, no warm-up steps, training state untouched)."""
        self.lr = float(lr)
        if momentum is not None:
            self.momentum = float(momentum)
        if weight_decay is not None:
            self.weight_decay = float(weight_decay)
        if self.buckets is not None:
            self.buckets.set_hyper(self.lr, self.momentum, self.weight_decay, self.nesterov)
        if self.graph is not None:
            self.graph = None
            self._recapture()

    def _recapture(self) -> None:
        torch.cuda.synchronize(self.device)
        graph = torch.cuda.CUDAGraph()
        steps_before, launches_before, native_before = self.sgd.steps, self.sgd.comm.launches, self._native_count()
        with torch.cuda.graph(graph):
            self._body()
        self.sgd.steps, self.sgd.comm.launches = steps_before, launches_before
        self._replayed_kernels -= self._native_count() - native_before
        torch.cuda.synchronize(self.device)
        self.graph = graph


class _LocalAlgo:
    """What TrainStep needs from an algorithm object when the step itself does no collective."""

    def __init__(self, comm):
        self.comm = comm
        self.steps = 0


class EATrainStep(TrainStep):
    """Elastic-averaging training on the flagship path: the reference's loop
    ``df(params, x, y)`` -> local SGD -> ``averageParameters(params)`` (examples/client_remote.lua:192-210,
    examples/mnist-ea.lua:100-110) with forward, backward and the local SGD kernel as ONE CUDA-graph replay per step,
    and every ``tau``-th step one more launch: the fused elastic-average kernel (csrc/fused_ea.cu, lua/AllReduceEA.lua:34-45).
    The center is double buffered, so the elastic kernel's arguments alternate and it stays outside the graph — one
    asynchronous launch every ``tau`` steps.  Nodes may take different numbers of steps: ``ea.synchronizeCenter()`` /
    ``ea.synchronizeParameters()`` drain as in the reference."""

    def __init__(self, model: torch.nn.Module, flat: FlatParams, ea, batch_shape: Tuple[int, ...], lr: float,
                 momentum: float = 0.0, weight_decay: float = 0.0, nesterov: bool = False, loss_fn: Optional[Callable] = None,
                 use_graph: bool = True, target_dtype: torch.dtype = torch.int64, warmup: int = 3):
        if momentum != 0.0 and flat.momentum is None:
            raise ValueError("momentum needs FlatParams(with_momentum=True)")
        super().__init__(model, flat, _LocalAlgo(ea.comm), batch_shape, lr, momentum, weight_decay, nesterov, loss_fn,
                         use_graph, "auto", target_dtype, warmup)
        self.ea = ea

    def _local_sgd(self, p_in=None) -> None:
        flat = self.flat
        n = flat.padded
        if self.device.type == "cuda":
            from ..ops.local import local_sgd
            local_sgd(flat.params.tensor[:n], flat.grads.tensor[:n], flat.momentum.tensor[:n] if self.momentum != 0.0 else None,
                      lr=self.lr, momentum=self.momentum, weight_decay=self.weight_decay, nesterov=self.nesterov,
                      zero_grads=True, p_in=p_in)
            self.sgd.comm.launches += 1
        else:
            with torch.no_grad():
                src = flat.params.tensor[:n] if p_in is None else p_in
                g = flat.grads.tensor[:n]
                if self.weight_decay:
                    g = g + self.weight_decay * src
                if self.momentum != 0.0:
                    m = flat.momentum.tensor[:n]
                    m.mul_(self.momentum).add_(g)
                    g = g + self.momentum * m if self.nesterov else m
                flat.params.tensor[:n].copy_(src - self.lr * g)
                flat.grads.tensor[:n].zero_()

    def _body(self) -> None:
        self._forward_backward()
        self._local_sgd()
        self.sgd.steps += 1

    def run(self) -> torch.Tensor:
        loss = super().run()
        self.ea.averageParameters(self.flat)      # counts the step; every tau-th call launches the fused elastic kernel
        return loss


class AsyncTrainStep(EATrainStep):
    """The asynchronous EASGD client on the flagship path.  The reference's client does
    ``grads at x`` -> ``syncClient`` (d = alpha (x - c); x -= d; c += d) -> ``x -= lr g`` (examples/EASGD_client.lua:106-117,
    lua/AsyncEA.lua:134-146).  d depends only on x and the center, not on the gradient, so on a sync step the exchange
    kernel (csrc/async_ea.cu) starts on a side stream at the SAME time as forward/backward, reads x, writes x - d into a
    second buffer and adds d into the server's center over NVLink; when backward is done the SGD kernel forms
    ``(x - d) - lr g``.  Same arithmetic and order of effects as the reference, with the exchange hidden under the compute.
    Two graphs are captured: the plain step and the sync step."""

    def __init__(self, model: torch.nn.Module, flat: FlatParams, client, batch_shape: Tuple[int, ...], lr: float, **kw):
        if client.role != "client":
            raise ValueError("AsyncTrainStep drives a client")
        super().__init__(model, flat, client, batch_shape, lr, **kw)
        self.client = client
        self.graph_sync: Optional[torch.cuda.CUDAGraph] = None
        self._moved = torch.empty_like(flat.params.tensor) if flat.device.type == "cuda" else None
        self._side = torch.cuda.Stream(flat.device) if flat.device.type == "cuda" else None

    def _sync_body(self) -> None:
        cl, flat = self.client, self.flat
        cur = torch.cuda.current_stream(self.device)
        self._side.wait_stream(cur)
        cl.comm.async_client_sync(cl.server_rank, flat.params.tensor, cl._center, cl._n, cl.alpha, cl.exclusive,
                                  stream=self._side, x_out=self._moved)
        self._forward_backward()
        cur.wait_stream(self._side)
        self._local_sgd(p_in=self._moved[:flat.padded])
        self.sgd.steps += 1

    def capture(self) -> None:
        if self._captured:
            return
        if self.client._center is None:
            raise RuntimeError("call client.initClient(flat) before the first step")
        super().capture()
        if self.graph is not None:
            steps_before, launches_before, native_before = self.sgd.steps, self.sgd.comm.launches, self._native_count()
            self.graph_sync = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph_sync):
                self._sync_body()
            self.sgd.steps, self.sgd.comm.launches = steps_before, launches_before
            self._replayed_kernels -= self._native_count() - native_before
            torch.cuda.synchronize(self.device)

    def run(self) -> torch.Tensor:
        if not self._captured:
            self.capture()
        cl = self.client
        cl.step += 1
        if cl.step % cl.tau != 0:
            return TrainStep.run(self)
        if self._staged:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(self._staged_ev)
            self.x.copy_(self._x_next, non_blocking=True)
            self.y.copy_(self._y_next, non_blocking=True)
            self._consumed_ev.record(cur)
            self._staged = False
        if self.graph_sync is not None:
            self.graph_sync.replay()
            self.sgd.steps += 1
            self.sgd.comm.launches += self.launches_per_step + 1
            self._replayed_kernels += self.kernels_per_step + 1
        elif self.device.type == "cuda":
            self._sync_body()
        else:
            self._forward_backward()
            cl.step -= 1
            cl.syncClient(self.flat)
            self._local_sgd()
            self.sgd.steps += 1
        return self.loss
