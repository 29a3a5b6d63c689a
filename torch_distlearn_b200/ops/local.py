"""Single-GPU flat-vector kernels (csrc/fused_sgd.cu: local_sgd, csrc/fused_ea.cu: local_ea,
csrc/local_ops.cu: sumsq_diff / confusion_add).  They raise if the CUDA library is missing —
there is no silent PyTorch fallback on a GPU machine."""
from __future__ import annotations

from typing import Optional

import torch

from .. import _lib


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def local_sgd(p: torch.Tensor, g: torch.Tensor, m: Optional[torch.Tensor] = None, lr: float = 0.01, momentum: float = 0.0,
              weight_decay: float = 0.0, nesterov: bool = False, grad_scale: float = 1.0, zero_grads: bool = False,
              p_in: Optional[torch.Tensor] = None) -> None:
    """p = p_in - lr * (scale*g [+ wd*p_in] [momentum]) on flat fp32 CUDA vectors (p_in defaults to p: in place);
    optional zero of g."""
    assert p.is_cuda and g.is_cuda and p.dtype == g.dtype == torch.float32 and p.is_contiguous() and g.is_contiguous()
    assert p.numel() == g.numel()
    src = p if p_in is None else p_in
    assert src.is_cuda and src.dtype == torch.float32 and src.is_contiguous() and src.numel() == p.numel()
    with torch.cuda.device(p.device):
        _lib.check(_lib.cuda().dl_local_sgd_from(p.data_ptr(), src.data_ptr(), g.data_ptr(), m.data_ptr() if m is not None else None,
                                                 p.numel(), lr, momentum, weight_decay, int(nesterov), grad_scale, int(zero_grads),
                                                 _stream(p)))


def local_ea(x: torch.Tensor, c: torch.Tensor, alpha: float) -> None:
    """Single-node elastic step: d = alpha (x - c); x -= d; c += d."""
    assert x.is_cuda and c.is_cuda and x.dtype == c.dtype == torch.float32 and x.numel() == c.numel()
    with torch.cuda.device(x.device):
        _lib.check(_lib.cuda().dl_local_ea(x.data_ptr(), c.data_ptr(), x.numel(), alpha, _stream(x)))


_scratch = {}


def sumsq_diff(a: torch.Tensor, b: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Deterministic sum((a-b)^2) (b=None: sum(a^2)) as a 1-element float64 CUDA tensor."""
    assert a.is_cuda and a.dtype == torch.float32 and a.is_contiguous()
    lib = _lib.cuda()
    key = (a.device, torch.cuda.current_stream(a.device).cuda_stream)
    sc = _scratch.get(key)
    if sc is None:
        sc = torch.zeros(int(lib.dl_sumsq_scratch_bytes()), dtype=torch.uint8, device=a.device)
        _scratch[key] = sc
    if out is None:
        out = torch.empty(1, dtype=torch.float64, device=a.device)
    with torch.cuda.device(a.device):
        _lib.check(lib.dl_sumsq_diff(a.data_ptr(), b.data_ptr() if b is not None else None, a.numel(), out.data_ptr(),
                                     sc.data_ptr(), _stream(a)))
    return out


def confusion_add(logits: torch.Tensor, target: torch.Tensor, mat: torch.Tensor) -> None:
    assert logits.is_cuda and logits.dtype == torch.float32 and logits.is_contiguous() and mat.dtype == torch.int64
    t = target.to(torch.int64).contiguous()
    with torch.cuda.device(logits.device):
        _lib.check(_lib.cuda().dl_confusion_add(logits.data_ptr(), t.data_ptr(), logits.shape[0], logits.shape[1],
                                                mat.data_ptr(), _stream(logits)))
