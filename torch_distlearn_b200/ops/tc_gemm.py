"""fp32-accurate GEMM on the Blackwell tensor cores (csrc/tc_gemm.cu: tcgen05.mma kind::tf32 with the
3xTF32 split, TMA operand loads, accumulator in TMEM, fp32 register flush every 128 K)."""
from __future__ import annotations

from typing import Optional

import torch

from .. import _lib


def tc_gemm(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None,
            split_k: int = 1, accumulate: bool = False, terms: int = 3, a_mn: bool = False, b_mn: bool = False) -> torch.Tensor:
    """``out[M,N] (+)= A @ B^T (+ bias[N])`` for 2-D fp32 CUDA tensors with a contiguous last dimension
    (row pitches may exceed the row length; multiples of 4 elements).

    ``a`` is ``[M, K]`` (K contiguous) — or, with ``a_mn=True``, the transposed storage ``[K, M]``;
    ``b`` is ``[N, K]`` — or, with ``b_mn=True``, ``[K, N]``.  ``terms=3`` gives fp32-level accuracy
    (3xTF32), ``terms=1`` plain TF32.  ``split_k`` splits the reduction over CTAs (partial tiles are
    added with atomics); ``accumulate`` adds into ``out`` instead of overwriting it."""
    assert a.is_cuda and b.is_cuda and a.dtype == b.dtype == torch.float32 and a.dim() == b.dim() == 2
    assert a.stride(1) == 1 and b.stride(1) == 1, "the last dimension of a and b must be contiguous"
    (M, K) = (a.shape[1], a.shape[0]) if a_mn else a.shape
    (N, Kb) = (b.shape[1], b.shape[0]) if b_mn else b.shape
    assert K == Kb, f"reduction lengths differ: {K} vs {Kb}"
    if out is None:
        assert not accumulate
        out = torch.empty(M, N, dtype=torch.float32, device=a.device)
    assert out.shape == (M, N) and out.stride(1) == 1 and out.dtype == torch.float32
    with torch.cuda.device(a.device):
        _lib.check(_lib.cuda().dl_tc_gemm_ex(a.data_ptr(), b.data_ptr(), out.data_ptr(),
                                             bias.data_ptr() if bias is not None else None, M, N, K, a.stride(0), b.stride(0),
                                             out.stride(0), int(a_mn), int(b_mn), int(split_k), int(accumulate), int(terms),
                                             torch.cuda.current_stream(a.device).cuda_stream))
    return out
