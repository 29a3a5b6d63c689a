"""Thin Python wrappers over the local (single-GPU) kernels of libdistlearn_b200."""
from .bn import FusedBNReLUPool
from .head import head_nll
from .conv import TCConv5x5, col2im5x5, conv5x5_nhwc, im2col5x5
from .local import confusion_add, local_ea, local_sgd, sumsq_diff
from .tc_gemm import tc_gemm

__all__ = ["FusedBNReLUPool", "TCConv5x5", "col2im5x5", "confusion_add", "conv5x5_nhwc", "head_nll", "im2col5x5", "local_ea", "local_sgd",
           "sumsq_diff", "tc_gemm"]
