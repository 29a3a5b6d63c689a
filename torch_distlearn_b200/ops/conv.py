"""5x5 / stride 1 / pad 2 convolution on the Blackwell tensor cores, forward and backward.

The reference's convolution is ``SpatialConvolutionMM`` (examples/Model.lua:20,26,32,38): im2col + SGEMM
in strict fp32.  Same decomposition here, NHWC activations, with the three GEMMs of a training step on
tcgen05 (csrc/tc_gemm.cu, 3xTF32 = fp32-level accuracy):

    forward   y[M, Cout]      = col[M, 25 Cin] @ W[Cout, 25 Cin]^T + bias          (K-major x K-major)
    wgrad     dW[Cout, 25Cin] = dY[M, Cout]^T @ col[M, 25 Cin]                     (MN-major x MN-major, split-K over M)
    dgrad     dcol[M, 25 Cin] = dY[M, Cout] @ W[Cout, 25 Cin]  ->  col2im -> dX    (K-major x MN-major)

with M = B*H*W and the (kh, kw, cin) column order of ``col``; no operand is ever transposed in memory.

When Cin and Cout are multiples of 32 (every layer of the CIFAR net but the first) even ``col`` does not
exist: the GEMM's TMA producer reads the activation tensor itself with 4-D boxes shifted by the filter
tap, out-of-image elements arriving as zeros (``dl_tc_conv5x5``: implicit GEMM for forward, dgrad and
wgrad).  The explicit im2col / col2im path remains for Cin = 3.
"""
from __future__ import annotations

import math
import os
from typing import Optional

import torch

from .. import _lib
from .tc_gemm import tc_gemm

TAPS = 25


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def im2col5x5(x: torch.Tensor, ld: Optional[int] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x [B,H,W,C] (contiguous NHWC) -> col [B*H*W, ld], ld = 25*C rounded up to a multiple of 4."""
    assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous() and x.dim() == 4
    B, H, W, C = x.shape
    ld = ld or (TAPS * C + 3) // 4 * 4
    if out is None:
        out = torch.empty(B * H * W, ld, dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.cuda().dl_im2col5x5(x.data_ptr(), out.data_ptr(), B, H, W, C, ld, _stream(x)))
    return out


def col2im5x5(dcol: torch.Tensor, shape) -> torch.Tensor:
    """dcol [B*H*W, ld] -> dx [B,H,W,C]."""
    B, H, W, C = shape
    assert dcol.is_cuda and dcol.is_contiguous() and dcol.shape[0] == B * H * W
    dx = torch.empty(B, H, W, C, dtype=torch.float32, device=dcol.device)
    with torch.cuda.device(dcol.device):
        _lib.check(_lib.cuda().dl_col2im5x5(dcol.data_ptr(), dx.data_ptr(), B, H, W, C, dcol.shape[1], _stream(dcol)))
    return dx


def _splits(m_tiles: int, n_tiles: int, k: int, sms: int = 148, overhead_kb: int = 6) -> int:
    """K-split that minimises  waves x (K-blocks per CTA + fixed per-CTA cost)  for a one-CTA-per-SM kernel:
    few output tiles need many splits to fill the 148 SMs, and the number of CTAs should land just under a
    whole number of waves.  Every split keeps at least 4 K-blocks (K = 128)."""
    tiles = max(m_tiles * n_tiles, 1)
    kb = max((k + 31) // 32, 1)
    best, best_cost = 1, None
    for s in range(1, max(1, kb // 4) + 1):
        per = (kb + s - 1) // s
        ctas = tiles * ((kb + per - 1) // per)
        cost = ((ctas + sms - 1) // sms) * (per + overhead_kb)
        if best_cost is None or cost < best_cost:
            best, best_cost = s, cost
    return best


def _implicit_ok(B: int, H: int, W: int, cin: int, cout: int) -> bool:
    """Geometry the 4-D TMA boxes support: whole image rows tile the 128- and 32-pixel boxes."""
    if cin % 32 or cout % 32 or W > 32 or 32 % W or 128 % W:
        return False
    for px in (128, 32):
        rows = px // W
        bh = min(rows, H)
        if H % bh or rows % bh:
            return False
    return (B * H * W) % 32 == 0


def _tc_conv(mode: int, act: torch.Tensor, w: torch.Tensor, out: torch.Tensor, bias, B, H, W, cin, cout, split_k, terms,
             accumulate: bool = False, w_split: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``w_split``: [2, Cout, 25*Cin] = (hi, lo) of ``w`` from :func:`split_weights` — modes 1 and 2 then take the weights
    already split (csrc/tc_gemm.cu: dl_tc_conv5x5_ws)."""
    with torch.cuda.device(act.device):
        _lib.check(_lib.cuda().dl_tc_conv5x5_ws(mode, act.data_ptr(), w.data_ptr(),
                                                w_split[0].data_ptr() if w_split is not None else None,
                                                w_split[1].data_ptr() if w_split is not None else None, out.data_ptr(),
                                                bias.data_ptr() if bias is not None else None, B, H, W, cin, cout, int(split_k),
                                                int(accumulate), int(terms), _stream(act)))
    return out


# Pre-split weights (hi = TF32(w), lo = w - hi), one persistent buffer per weight tensor: made once per forward pass and used
# by the forward and the dgrad GEMM of that step, instead of being re-made by the converter warps of every CTA in every K
# block.  OFF by default: measured on a B200 it is correct (tests/test_conv_presplit_gpu.py) and changes nothing — the
# forward GEMMs take 71.0 / 67.1 us with it against 71.2 / 67.3 without, the step 1.141 ms against 1.121 with the three
# split launches added (profiles/r2/r2k_*) — which is what says the GEMM is not bound by its converter warps
# (profiles/ncu_gemm_r2.md).  DL_B200_TC_PRESPLIT=1 or presplit(True) switches it on.
_PRESPLIT = os.environ.get("DL_B200_TC_PRESPLIT", "0") == "1"
_SPLITS: dict = {}


def presplit(on: Optional[bool] = None) -> bool:
    """Query / set whether the convolutions hand their weights to the GEMM already split."""
    global _PRESPLIT
    before = _PRESPLIT
    if on is not None:
        _PRESPLIT = bool(on)
    return before


class _Split:
    __slots__ = ("buf", "seq")

    def __init__(self, w: torch.Tensor):
        self.buf = torch.empty((2,) + tuple(w.shape), dtype=torch.float32, device=w.device)
        self.seq = 0


def split_weights(w: torch.Tensor) -> "_Split":
    """(Re)compute the split of ``w`` into its persistent buffer on the current stream; returns the cache entry
    (``.buf`` [2, ...], ``.seq`` counts the recomputations so a backward pass can tell whether the buffer still is its own)."""
    key = (w.data_ptr(), tuple(w.shape), w.device.index)
    e = _SPLITS.get(key)
    if e is None:
        if len(_SPLITS) >= 64:      # short-lived weight tensors (tests, sweeps): do not let their buffers pile up
            _SPLITS.clear()         # (a pending backward pass keeps its own entry alive through ctx.split)
        e = _SPLITS[key] = _Split(w)
    with torch.cuda.device(w.device):
        _lib.check(_lib.cuda().dl_split_tf32(w.data_ptr(), e.buf[0].data_ptr(), e.buf[1].data_ptr(), w.numel(), _stream(w)))
    e.seq += 1
    return e


class _Conv5x5Implicit(torch.autograd.Function):
    """Forward / dgrad / wgrad as implicit GEMMs: no im2col, col2im or transposed buffer at all."""

    @staticmethod
    def forward(ctx, x, weight, bias, terms):
        x = x.contiguous()
        B, H, W, C = x.shape
        cout = weight.shape[0]
        M = B * H * W
        bn = 64 if cout <= 64 else 128
        y = torch.empty(M, cout, dtype=torch.float32, device=x.device)
        split = split_weights(weight.detach()) if (_PRESPLIT and terms == 3 and weight.is_contiguous()) else None
        _tc_conv(1, x, weight, y, bias, B, H, W, C, cout, _splits((M + 127) // 128, (cout + bn - 1) // bn, TAPS * C), terms,
                 w_split=split.buf if split is not None else None)
        ctx.save_for_backward(x, weight)
        ctx.split = (split, split.seq) if split is not None else None
        ctx.meta = (terms, bias is not None)
        ctx.params = (weight, bias)
        return y.view(B, H, W, cout)

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        terms, has_bias = ctx.meta
        B, H, W, C = x.shape
        cout, kc = weight.shape
        M = B * H * W
        dy = dy.contiguous()
        dx = dw = db = None
        from .bn import colsum, grad_done, grad_sink
        if ctx.needs_input_grad[1]:
            sink = grad_sink(ctx.params[0])
            if sink is not None:       # the wgrad GEMM's epilogue adds straight into weight.grad (red.add): no extra pass
                _tc_conv(3, x, dy, sink, None, B, H, W, C, cout, _splits((cout + 127) // 128, (kc + 127) // 128, M), terms,
                         accumulate=True)
            else:
                dw = torch.empty(cout, kc, dtype=torch.float32, device=x.device)
                _tc_conv(3, x, dy, dw, None, B, H, W, C, cout, _splits((cout + 127) // 128, (kc + 127) // 128, M), terms)
        bias_done = False
        if has_bias and ctx.needs_input_grad[2]:
            from .bn import take_bias_done
            bias_done = take_bias_done(ctx.params[1])      # the BatchNorm block behind us already added it into bias.grad
            if not bias_done:
                db = colsum(dy.view(M, cout), out=grad_sink(ctx.params[1]))
        if ctx.needs_input_grad[0]:
            bn = 64 if C <= 64 else 128
            dx = torch.empty(B, H, W, C, dtype=torch.float32, device=x.device)
            w_split = None
            if ctx.split is not None:
                split, seq = ctx.split
                if split.seq != seq:          # another forward pass has used the buffer since: make it ours again
                    split = split_weights(weight.detach())
                w_split = split.buf
            _tc_conv(2, dy, weight, dx, None, B, H, W, C, cout, _splits((M + 127) // 128, (C + bn - 1) // bn, TAPS * cout), terms,
                     w_split=w_split)
        # only now, with the dgrad GEMM (which reads the weights) enqueued, may a listener start updating this layer
        for p, g in zip(ctx.params, (dw, db)):
            if p is not None and g is None:
                grad_done(p)
        return dx, dw, db, None


class _Conv5x5(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, terms):
        x = x.contiguous()
        B, H, W, C = x.shape
        cout, kc = weight.shape
        assert kc == TAPS * C, f"weight must be [Cout, 25*Cin] = [{cout}, {TAPS * C}]"
        col = im2col5x5(x)
        ld = col.shape[1]
        w = weight
        if ld != kc:                                    # Cin = 3: pad K to a multiple of 4 with zero columns
            w = torch.zeros(cout, ld, dtype=weight.dtype, device=weight.device)
            w[:, :kc] = weight
        M = B * H * W
        bn = 64 if cout <= 64 else 128
        y = tc_gemm(col, w, bias=bias, split_k=_splits((M + 127) // 128, (cout + bn - 1) // bn, ld), terms=terms)
        ctx.save_for_backward(col, w)
        ctx.meta = (B, H, W, C, cout, kc, terms, bias is not None)
        ctx.params = (weight, bias)
        return y.view(B, H, W, cout)

    @staticmethod
    def backward(ctx, dy):
        col, w = ctx.saved_tensors
        B, H, W, C, cout, kc, terms, has_bias = ctx.meta
        M, ld = col.shape
        dy2 = dy.contiguous().view(M, cout)
        dx = dw = db = None
        if ctx.needs_input_grad[1]:
            bn = 64 if ld <= 64 else 128
            dw_full = tc_gemm(dy2, col, a_mn=True, b_mn=True, terms=terms,
                              split_k=_splits((cout + 127) // 128, (ld + bn - 1) // bn, M))
            dw = dw_full[:, :kc] if ld != kc else dw_full
        bias_done = False
        if has_bias and ctx.needs_input_grad[2]:
            from .bn import colsum, grad_done, grad_sink, take_bias_done
            bias_done = take_bias_done(ctx.params[1])
            if not bias_done:
                db = colsum(dy2, out=grad_sink(ctx.params[1]))
        if ctx.needs_input_grad[0]:
            dcol = tc_gemm(dy2, w, b_mn=True, terms=terms)          # [M, ld]; K = Cout is short, no split
            dx = col2im5x5(dcol, (B, H, W, C))
        if has_bias and ctx.needs_input_grad[2] and db is None:
            grad_done(ctx.params[1])
        return dx, dw, db, None


def conv5x5_nhwc(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, terms: int = 3,
                  implicit: Optional[bool] = None) -> torch.Tensor:
    """x [B,H,W,Cin] NHWC, weight [Cout, 25*Cin] with columns ordered (kh, kw, cin) -> y [B,H,W,Cout]."""
    B, H, W, C = x.shape
    if implicit is None:
        implicit = _implicit_ok(B, H, W, C, weight.shape[0])
    return (_Conv5x5Implicit if implicit else _Conv5x5).apply(x, weight, bias, terms)


class TCConv5x5(torch.nn.Module):
    """``nn.Conv2d(cin, cout, 5, 1, 2)`` with NHWC activations and the weight kept as the GEMM operand
    ``[cout, 25*cin]`` ((kh, kw, cin) column order).  Same initialisation distribution as nn.Conv2d."""

    def __init__(self, cin: int, cout: int, terms: int = 3, implicit: Optional[bool] = None):
        super().__init__()
        self.cin, self.cout, self.terms, self.implicit = cin, cout, terms, implicit
        self.weight = torch.nn.Parameter(torch.empty(cout, TAPS * cin))
        self.bias = torch.nn.Parameter(torch.empty(cout))
        bound = 1.0 / math.sqrt(cin * TAPS)
        torch.nn.init.uniform_(self.weight, -bound, bound)       # kaiming_uniform(a=sqrt(5)) == U(-1/sqrt(fan_in), +)
        torch.nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return conv5x5_nhwc(x, self.weight, self.bias, self.terms, self.implicit)

    @torch.no_grad()
    def load_conv2d(self, conv: torch.nn.Conv2d) -> None:
        """Copy the weights of an ``nn.Conv2d`` ([cout, cin, 5, 5] -> [cout, (kh, kw, cin)])."""
        self.weight.copy_(conv.weight.permute(0, 2, 3, 1).reshape(self.cout, -1))
        self.bias.copy_(conv.bias)

    def extra_repr(self) -> str:
        return f"{self.cin}, {self.cout}, kernel_size=5, padding=2, layout=NHWC, gemm=tcgen05 {'3xTF32' if self.terms == 3 else 'TF32'}"
