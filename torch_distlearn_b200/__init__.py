"""torch_distlearn_b200 — a Blackwell-native distributed-learning library with the capabilities
of DistLearn (shanlior/torch-distlearn): AllReduceSGD, AllReduceEA and AsyncEA over flat
parameter/gradient vectors in a symmetric heap, driven by hand-written sm_100a kernels that read
and write peer GPUs directly over NVLink 5 / NVSwitch.

    import torch_distlearn_b200 as dl
    comm  = dl.P2PCommunicator()                    # one process per GPU (torchrun)
    flat  = dl.FlatParams(comm, model)              # params / grads become views of symmetric buffers
    sgd   = dl.AllReduceSGD(comm, flat)
    sgd.synchronizeParameters()
    loss.backward(); sgd.step(lr=0.1)               # ONE fused kernel: all-reduce + 1/n + SGD + all-gather
"""
from .parallel.comm import (CommResult, Communicator, DistCommunicator, EmulatedCommunicator, HostCommunicator, P2PCommunicator,
                            SymBuffer, create_communicator)
from .parallel.flat import FlatParams, TableStager, walk_table
from .parallel.allreduce_sgd import AllReduceSGD
from .parallel.allreduce_ea import AllReduceEA
from .parallel.async_ea import AsyncEA
from .parallel.tree import LocalhostTree, Tree
from .parallel.trainer import TrainStep
from .parallel.bucketed import GradBuckets
from .utils.checkpoint import load_checkpoint, save_checkpoint
from .utils.color_print import print_client, print_server

__version__ = "0.1.0"

__all__ = [
    "CommResult", "Communicator", "DistCommunicator", "EmulatedCommunicator", "HostCommunicator", "P2PCommunicator", "SymBuffer",
    "create_communicator", "FlatParams", "TableStager", "walk_table", "AllReduceSGD", "AllReduceEA", "AsyncEA",
    "LocalhostTree", "Tree", "TrainStep", "GradBuckets", "make_communicator", "save_checkpoint", "load_checkpoint", "print_client", "print_server",
]


def __getattr__(name):
    # lazy: importing .launch here would shadow `python -m torch_distlearn_b200.launch`
    if name == "make_communicator":
        from .launch import make_communicator
        return make_communicator
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
