"""Fused training-mode BatchNorm + ReLU + 2x2 max-pool on NHWC activations (csrc/bn_ops.cu): the
elementwise tail of each block of the reference's CIFAR network (examples/Model.lua:21-24) in two
passes over the activation forward and two backward, instead of one pass per op."""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .. import _lib


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def grad_sink(p) -> "torch.Tensor | None":
    """The tensor a backward kernel may accumulate a parameter gradient into directly, or None.

    Autograd adds a returned gradient into ``p.grad`` with one extra element-wise kernel per parameter; the
    backward kernels of this package can add into ``p.grad`` themselves (it is a view of the flat gradient vector,
    zeroed by the fused SGD kernel).  That is only done when it is indistinguishable from autograd's accumulation:
    ``p.grad`` exists and is dense fp32 of the parameter's shape, and no tensor hook wants to see (or change) the
    incoming gradient.  Whoever writes into the sink must call :func:`grad_done` afterwards."""
    if not DIRECT_GRAD or not isinstance(p, torch.nn.Parameter):
        return None
    g = p.grad
    if g is None or g.dtype != torch.float32 or g.shape != p.shape or not g.is_contiguous() or not g.is_cuda:
        return None
    if getattr(p, "_backward_hooks", None):
        return None
    return g


def grad_done(p) -> None:
    """Run the parameter's post-accumulate-grad hooks (e.g. GradBuckets' "this bucket is ready") after a backward
    kernel has been enqueued that adds into ``p.grad`` directly — autograd would have run them after its own add."""
    hooks = getattr(p, "_post_accumulate_grad_hooks", None)
    if hooks:
        for h in list(hooks.values()):
            h(p)


DIRECT_GRAD = True

# Bias parameters whose gradient was already produced by the BatchNorm block behind their convolution (the column sums of
# dy come out of the block's backward pass for free); the convolution's backward finds its bias here and skips its own
# reduction over dy.  Filled and emptied within one backward pass.
_BIAS_DONE = set()


def take_bias_done(p) -> bool:
    if id(p) in _BIAS_DONE:
        _BIAS_DONE.discard(id(p))
        return True
    return False


def colsum(x: torch.Tensor, out: "torch.Tensor | None" = None) -> "torch.Tensor | None":
    """Column sums of a contiguous [M, C] fp32 CUDA matrix (C % 4 == 0): the bias gradient of a convolution.
    With ``out`` the sums are ADDED into it and None is returned."""
    M, C = x.shape
    if not (x.is_cuda and x.is_contiguous() and x.dtype == torch.float32 and C % 4 == 0 and C <= 1024):
        if out is None:
            return x.sum(0)
        out.add_(x.sum(0))
        return None
    acc = out is not None
    if out is None:
        out = torch.empty(C, dtype=torch.float32, device=x.device)
    work = torch.empty(int(_lib.cuda().dl_bn_work_bytes(C)) // 8 + 1, dtype=torch.float64, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.cuda().dl_colsum(x.data_ptr(), out.data_ptr(), M, C, int(acc), work.data_ptr(), _stream(x)))
    return None if acc else out


class _BNReLUPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y, gamma, beta, running_mean, running_var, eps, momentum, conv_bias=None):
        ctx.conv_bias = conv_bias
        y = y.contiguous()
        B, H, W, C = y.shape
        p = torch.empty(B, H // 2, W // 2, C, dtype=torch.float32, device=y.device)
        mean = torch.empty(C, dtype=torch.float32, device=y.device)
        invstd = torch.empty(C, dtype=torch.float32, device=y.device)
        work = torch.empty(int(_lib.cuda().dl_bn_work_bytes(C)) // 8 + 1, dtype=torch.float64, device=y.device)
        with torch.cuda.device(y.device):
            _lib.check(_lib.cuda().dl_bn_relu_pool_fwd(
                y.data_ptr(), p.data_ptr(), gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                running_mean.data_ptr() if running_mean is not None else None,
                running_var.data_ptr() if running_var is not None else None, B, H, W, C, float(eps), float(momentum),
                work.data_ptr(), _stream(y)))
        ctx.save_for_backward(y, gamma, beta, mean, invstd)
        ctx.params = (gamma, beta)
        return p

    @staticmethod
    def backward(ctx, dp):
        y, gamma, beta, mean, invstd = ctx.saved_tensors
        B, H, W, C = y.shape
        dp = dp.contiguous()
        dy = torch.empty_like(y)
        sink_g, sink_b = grad_sink(ctx.params[0]), grad_sink(ctx.params[1])
        direct = sink_g is not None and sink_b is not None
        dgamma = sink_g if direct else torch.empty(C, dtype=torch.float32, device=y.device)
        dbeta = sink_b if direct else torch.empty(C, dtype=torch.float32, device=y.device)
        work = torch.empty(int(_lib.cuda().dl_bn_work_bytes(C)) // 8 + 1, dtype=torch.float64, device=y.device)
        # the convolution's bias gradient = column sums of dy: produced here, in the pass that writes dy, when the bias can
        # take it directly (same conditions as for our own parameters)
        sink_cb = grad_sink(ctx.conv_bias) if (direct and ctx.conv_bias is not None and ctx.conv_bias.requires_grad) else None
        with torch.cuda.device(y.device):
            _lib.check(_lib.cuda().dl_bn_relu_pool_bwd_bias(
                y.data_ptr(), dp.data_ptr(), dy.data_ptr(), gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                dgamma.data_ptr(), dbeta.data_ptr(), sink_cb.data_ptr() if sink_cb is not None else None, B, H, W, C, int(direct),
                work.data_ptr(), _stream(y)))
        if sink_cb is not None:
            _BIAS_DONE.add(id(ctx.conv_bias))
        if direct:
            grad_done(ctx.params[0])
            grad_done(ctx.params[1])
            return dy, None, None, None, None, None, None, None
        return dy, dgamma, dbeta, None, None, None, None, None


class FusedBNReLUPool(torch.nn.Module):
    """``max_pool2d(relu(BatchNorm2d(C, eps)(x)), 2, 2)`` on NHWC tensors ``[B, H, W, C] -> [B, H/2, W/2, C]``.
    Parameters and buffers carry the names of ``nn.BatchNorm2d`` so state dicts are interchangeable."""

    def __init__(self, channels: int, eps: float = 1e-3, momentum: float = 0.1):
        super().__init__()
        self.channels, self.eps, self.momentum = channels, eps, momentum
        self.weight = torch.nn.Parameter(torch.ones(channels))
        self.bias = torch.nn.Parameter(torch.zeros(channels))
        self.register_buffer("running_mean", torch.zeros(channels))
        self.register_buffer("running_var", torch.ones(channels))
        self.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long))
        self._conv_bias = None

    def attach_conv_bias(self, bias: torch.nn.Parameter) -> None:
        """``bias`` belongs to the convolution whose output this block normalises: its gradient (the column sums of the
        gradient this block's backward produces) is then computed inside that backward pass instead of by a separate
        reduction in the convolution's backward.  Not a module attribute on purpose: the parameter stays the convolution's."""
        object.__setattr__(self, "_conv_bias", bias)

    def forward(self, y: torch.Tensor) -> torch.Tensor:
        if self.training and y.is_cuda:
            # (num_batches_tracked is not advanced: it only matters for momentum=None, and a host-side
            # counter update would be one more tiny kernel per layer per step)
            return _BNReLUPool.apply(y, self.weight, self.bias, self.running_mean, self.running_var, self.eps, self.momentum,
                                     self._conv_bias)
        z = F.batch_norm(y.permute(0, 3, 1, 2), self.running_mean, self.running_var, self.weight, self.bias, self.training,
                         self.momentum, self.eps)
        return F.max_pool2d(F.relu(z), 2, 2).permute(0, 2, 3, 1)

    def extra_repr(self) -> str:
        return f"{self.channels}, eps={self.eps}, momentum={self.momentum}, fused=BN+ReLU+MaxPool2x2 (NHWC)"
