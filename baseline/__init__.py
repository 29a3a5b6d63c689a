"""The measured baseline: torch.distributed (NCCL on GPUs, gloo on CPU)
implementations of the same three algorithms.  A path that only calls NCCL is the baseline, never
the product — nothing in torch_distlearn_b200 imports this package."""
