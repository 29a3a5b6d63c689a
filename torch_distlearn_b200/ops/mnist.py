"""The reference's MNIST network (examples/mnist.lua:53-81) trained by ONE kernel launch per step (csrc/mnist_step.cu:
an 8-CTA thread-block cluster, channels split over the CTAs, activations exchanged through distributed shared memory)."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from .. import _lib

FUSED_MAX_BATCH = 8
PARAM_NAMES = ("conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "linear.weight", "linear.bias")
PARAM_SHAPES = ((16, 1, 5, 5), (16,), (16, 16, 5, 5), (16,), (10, 400), (10,))


def _ptrs(tensors: Sequence[torch.Tensor]):
    arr = (C.c_void_p * 6)()
    for i, (t, shape) in enumerate(zip(tensors, PARAM_SHAPES)):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and tuple(t.shape) == shape):
            raise ValueError(f"mnist_step: tensor {i} must be a contiguous fp32 CUDA tensor of shape {shape}, got {tuple(t.shape)}")
        arr[i] = t.data_ptr()
    return arr


def mnist_step(x: torch.Tensor, y: torch.Tensor, params: Sequence[torch.Tensor], grads: Sequence[torch.Tensor], loss: torch.Tensor,
               logp: Optional[torch.Tensor] = None, lr: float = 0.0, apply_sgd: bool = False) -> None:
    """Forward + backward (+ ``p -= lr * g`` when ``apply_sgd`` and the batch is at most ``FUSED_MAX_BATCH``) of the MNIST
    convnet on ``x [B, 1024]``, ``y [B]`` (int64).  Without the fused update the gradient of the mean loss is ADDED into
    ``grads``.  ``loss``: 1-element fp32 CUDA tensor (overwritten), ``logp``: optional [B, 10] output."""
    B = x.shape[0]
    assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous() and x.numel() == B * 1024
    assert y.is_cuda and y.dtype == torch.int64 and y.is_contiguous() and y.numel() == B
    assert loss.is_cuda and loss.dtype == torch.float32
    if logp is not None:
        assert logp.is_cuda and logp.dtype == torch.float32 and logp.is_contiguous() and logp.numel() == B * 10
    with torch.cuda.device(x.device):
        _lib.check(_lib.cuda().dl_mnist_step(x.data_ptr(), y.data_ptr(), _ptrs(params), _ptrs(grads), loss.data_ptr(),
                                             logp.data_ptr() if logp is not None else None, B, float(lr), int(apply_sgd),
                                             torch.cuda.current_stream(x.device).cuda_stream))


def model_tensors(model: torch.nn.Module):
    """(params, grads) of a ``MnistConvNet`` in the kernel's order; grads must exist (FlatParams attaches them)."""
    named = dict(model.named_parameters())
    params = [named[n].data for n in PARAM_NAMES]
    grads = [named[n].grad for n in PARAM_NAMES]
    if any(g is None for g in grads):
        raise ValueError("mnist_step needs allocated gradients (FlatParams, or zero them once)")
    return params, grads
