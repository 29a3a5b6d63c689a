"""Data-parallel algorithms and the transport under them (the reference implements
synchronous gradient all-reduce, synchronous elastic averaging and an asynchronous center server —
data parallelism only; there is no TP/PP/SP/EP in it and none is invented here)."""
from .comm import (CommResult, Communicator, EmulatedCommunicator, HostCommunicator, P2PCommunicator, SymBuffer,
                   create_communicator)
from .flat import FlatParams, TableStager, walk_table
from .allreduce_sgd import AllReduceSGD
from .allreduce_ea import AllReduceEA
from .async_ea import AsyncEA
from .tree import LocalhostTree, Tree
