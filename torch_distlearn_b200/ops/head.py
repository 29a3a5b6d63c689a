"""Fused classifier head (csrc/head_ops.cu): Reshape -> Linear -> LogSoftMax -> ClassNLLCriterion of the
reference's CIFAR network (examples/Model.lua:44-52) on the NHWC feature map, two launches forward and two
backward instead of ~14 small torch kernels."""
from __future__ import annotations

import torch

from .. import _lib
from .bn import grad_done, grad_sink


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


class _HeadNLL(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feat, weight, bias, target):
        feat = feat.contiguous()
        B, H, W, C = feat.shape
        J = weight.shape[0]
        logp = torch.empty(B, J, dtype=torch.float32, device=feat.device)
        loss = torch.empty((), dtype=torch.float32, device=feat.device)
        rowloss = torch.empty(B, dtype=torch.float32, device=feat.device)
        tgt = target.to(torch.int64).contiguous()
        with torch.cuda.device(feat.device):
            _lib.check(_lib.cuda().dl_head_fwd(feat.data_ptr(), weight.data_ptr(), bias.data_ptr(), tgt.data_ptr(), logp.data_ptr(),
                                               loss.data_ptr(), rowloss.data_ptr(), B, H * W, C, J, _stream(feat)))
        ctx.save_for_backward(feat, weight, logp, tgt)
        ctx.params = (weight, bias)
        ctx.mark_non_differentiable(logp)
        return loss, logp

    @staticmethod
    def backward(ctx, gloss, _glogp):
        feat, weight, logp, tgt = ctx.saved_tensors
        B, H, W, C = feat.shape
        J = weight.shape[0]
        gloss = gloss.contiguous().to(torch.float32)
        dfeat = torch.empty_like(feat)
        sw, sb = grad_sink(ctx.params[0]), grad_sink(ctx.params[1])
        direct = sw is not None and sb is not None
        dw = sw if direct else torch.empty_like(weight)
        db = sb if direct else torch.empty(J, dtype=torch.float32, device=feat.device)
        with torch.cuda.device(feat.device):
            _lib.check(_lib.cuda().dl_head_bwd(feat.data_ptr(), weight.data_ptr(), logp.data_ptr(), tgt.data_ptr(), gloss.data_ptr(),
                                               dfeat.data_ptr(), dw.data_ptr(), db.data_ptr(), B, H * W, C, J, int(direct),
                                               _stream(feat)))
        if direct:
            grad_done(ctx.params[0])
            grad_done(ctx.params[1])
            return dfeat, None, None, None
        return dfeat, dw, db, None


def head_nll(feat: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, target: torch.Tensor):
    """feat [B,H,W,C] NHWC, weight [J, C*H*W] with the reference's (c, h, w) column order, bias [J], target [B]
    -> (mean NLL loss, log-probabilities [B, J]); J must be 10 or 16, fp32 CUDA tensors."""
    return _HeadNLL.apply(feat, weight, bias, target)


def head_supported(feat: torch.Tensor, weight: torch.Tensor) -> bool:
    return (feat.is_cuda and feat.dtype == torch.float32 and weight.is_contiguous() and weight.shape[0] in (10, 16)
            and feat.shape[0] * weight.shape[0] * 4 <= 48 * 1024)
