"""Thin Python wrappers over the local (single-GPU) kernels of libdistlearn_b200."""
from .local import confusion_add, local_ea, local_sgd, sumsq_diff

__all__ = ["confusion_add", "local_ea", "local_sgd", "sumsq_diff"]
