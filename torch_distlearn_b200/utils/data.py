"""Synthetic, partitioned datasets and batchers (reference: examples/Data.lua and the
``dataset.Dataset(...).sampledBatcher{...}`` calls in examples/mnist.lua:26-40,
examples/cifar10.lua:41-71).

The reference downloads MNIST from a URL and reads CIFAR-10 from hard-coded home directories;
neither exists here, so datasets are synthetic tensors of the same shapes (BASELINE.json).  To make
training curves meaningful the synthetic data is *learnable*: each class has a fixed random
prototype and samples are prototype + noise.

``Dataset(..., partition=i, partitions=N)`` gives node i (1-based) every N-th sample, so partitions
can have **uneven sizes** — the reason AllReduceSGD/EA carry the uneven-steps machinery
(lua/AllReduceSGD.lua:22).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Optional, Sequence, Tuple

import torch

CIFAR_CLASSES = ["airplane", "automobile", "bird", "cat", "deer", "dog", "frog", "horse", "ship", "truck"]
MNIST_CLASSES = [str(i) for i in range(10)]


@dataclass
class Batch:
    """What ``getBatch()`` returns in the reference: ``.input, .target, .batchSize``
    (examples/cifar10.lua:176-178,194)."""
    input: torch.Tensor
    target: torch.Tensor
    batchSize: int


class SyntheticDataset:
    """``num_samples`` items of shape ``input_dims`` in ``num_classes`` classes; this node sees the
    samples ``partition-1, partition-1+partitions, ...`` (1-based partition, as in the reference)."""

    def __init__(self, num_samples: int, input_dims: Sequence[int], num_classes: int = 10, partition: int = 1,
                 partitions: int = 1, seed: int = 1234, noise: float = 0.5):
        if not (1 <= partition <= partitions):
            raise ValueError("partition must be in 1..partitions")
        g = torch.Generator().manual_seed(seed)
        self.input_dims = tuple(input_dims)
        self.num_classes = num_classes
        self.prototypes = torch.randn(num_classes, *self.input_dims, generator=g)
        labels = torch.randint(0, num_classes, (num_samples,), generator=g)
        idx = torch.arange(partition - 1, num_samples, partitions)
        self.labels = labels[idx]
        self.noise = noise
        self.seed = seed * 7919 + partition
        self.partition, self.partitions = partition, partitions

    def __len__(self) -> int:
        return int(self.labels.numel())

    def materialize(self, indices: torch.Tensor, generator: torch.Generator) -> Tuple[torch.Tensor, torch.Tensor]:
        y = self.labels[indices]
        x = self.prototypes[y] + self.noise * torch.randn(len(indices), *self.input_dims, generator=generator)
        return x, y

    def sampledBatcher(self, batchSize: int, samplerKind: str = "permutation", cuda: bool = False,
                       device: Optional[torch.device] = None, pin: bool = True, inputDims=None, verbose: bool = False,
                       processor: Optional[Callable] = None):
        """Returns ``(getBatch, numBatches)`` like the reference's batcher.  Sampler kinds:
        ``permutation`` (examples/mnist.lua:32) and ``label-uniform`` (examples/cifar10.lua:54)."""
        if samplerKind not in ("permutation", "label-uniform"):
            raise ValueError(f"unknown samplerKind {samplerKind!r}")
        n = len(self)
        nb = (n + batchSize - 1) // batchSize if n else 0
        gen = torch.Generator().manual_seed(self.seed)
        dev = device if device is not None else (torch.device("cuda") if cuda else torch.device("cpu"))
        state = {"perm": None, "pos": 0}
        by_class = [torch.nonzero(self.labels == c).flatten() for c in range(self.num_classes)]
        nonempty = [c for c in range(self.num_classes) if len(by_class[c])]

        def next_indices() -> torch.Tensor:
            if samplerKind == "permutation":
                if state["perm"] is None or state["pos"] >= n:
                    state["perm"], state["pos"] = torch.randperm(n, generator=gen), 0
                idx = state["perm"][state["pos"]:state["pos"] + batchSize]
                state["pos"] += batchSize
                return idx
            # label-uniform: pick a class uniformly, then a sample of that class
            cls = torch.randint(0, len(nonempty), (batchSize,), generator=gen)
            out = torch.empty(batchSize, dtype=torch.long)
            for i, c in enumerate(cls.tolist()):
                pool = by_class[nonempty[c]]
                out[i] = pool[torch.randint(0, len(pool), (1,), generator=gen)]
            return out

        def getBatch() -> Batch:
            idx = next_indices()
            x, y = self.materialize(idx, gen)
            if processor is not None:
                x = processor(x)
            if dev.type == "cuda":
                if pin:
                    x, y = x.pin_memory(), y.pin_memory()
                x, y = x.to(dev, non_blocking=True), y.to(dev, non_blocking=True)
            return Batch(x, y, int(len(idx)))

        return getBatch, (lambda: nb)

    sampled_batcher = sampledBatcher

    # ---- native path: the partition materialised once as bytes, batches assembled by C++ threads ----------------
    U8_RANGE = 8.0      # samples are prototype (~N(0,1)) + noise: [-4, 4) covers them; quantisation step 1/32

    def materialize_u8(self) -> torch.Tensor:
        """Every sample of this partition as uint8 ``[n, *input_dims]`` (what a dataset on disk looks like):
        ``x = u8 * (U8_RANGE / 255) - U8_RANGE / 2`` gives the float sample back to within one quantisation step."""
        gen = torch.Generator().manual_seed(self.seed + 17)
        n = len(self)
        out = torch.empty((n,) + self.input_dims, dtype=torch.uint8)
        for lo in range(0, n, 4096):                      # bounded temporaries
            idx = torch.arange(lo, min(lo + 4096, n))
            x, _ = self.materialize(idx, gen)
            out[lo:lo + len(idx)] = ((x / self.U8_RANGE + 0.5) * 255).round().clamp_(0, 255).to(torch.uint8)
        return out

    def nativeBatcher(self, batchSize: int, samplerKind: str = "permutation", threads: int = 2, depth: int = 4, **kw):
        """A :class:`~torch_distlearn_b200.utils.native_loader.NativeBatcher` over this partition (csrc/loader.cpp): the
        Python batcher above costs ~5 ms per 128-sample CIFAR batch (`profiles/loader_cpu.jsonl`), several GPU steps."""
        from .native_loader import NativeBatcher
        if not hasattr(self, "_u8"):
            self._u8 = self.materialize_u8()
        return NativeBatcher(self._u8, self.labels, batchSize, sampler=samplerKind, num_classes=self.num_classes,
                             scale=self.U8_RANGE / 255, mean=[self.U8_RANGE / 2], std=[1.0], seed=self.seed, threads=threads,
                             depth=depth, **kw)


def mnist_dataset(partition: int = 1, partitions: int = 1, num_samples: int = 60000, seed: int = 1) -> SyntheticDataset:
    """MNIST-shaped: 1024-vectors (32x32 padded digits, examples/mnist.lua:34,53)."""
    return SyntheticDataset(num_samples, (1024,), 10, partition, partitions, seed)


def cifar_dataset(partition: int = 1, partitions: int = 1, num_samples: int = 50000, seed: int = 2) -> SyntheticDataset:
    """CIFAR-10-shaped: 3x32x32 (examples/cifar10.lua:56)."""
    return SyntheticDataset(num_samples, (3, 32, 32), 10, partition, partitions, seed)


class Data:
    """``Data(partitionNum, partitions, batchSize, cuda)`` -> getTrainingBatch, numTrainingBatches,
    getTestBatch, numTestBatches, classes  (examples/Data.lua:5-74)."""

    def __init__(self, partitionNum: int, partitions: int, batchSize: int, cuda: bool = False, train_samples: int = 50000,
                 test_samples: int = 10000, device: Optional[torch.device] = None, native: bool = True, threads: int = 2):
        self.train = cifar_dataset(partitionNum, partitions, train_samples, seed=2)
        self.test = cifar_dataset(partitionNum, partitions, test_samples, seed=3)
        self.test.prototypes = self.train.prototypes  # same distribution
        self.classes = list(CIFAR_CLASSES)
        self._native = []
        if native:
            # batches come from C++ worker threads, already normalised, in pinned memory (utils/native_loader.py)
            dev = device if device is not None else (torch.device("cuda") if cuda else torch.device("cpu"))
            tr = self.train.nativeBatcher(batchSize, "label-uniform", threads=threads)
            te = self.test.nativeBatcher(batchSize, "permutation", threads=1)
            self._native = [tr, te]
            self.getTrainingBatch, self.numTrainingBatches = tr.sampledBatcher(dev)
            self.getTestBatch, self.numTestBatches = te.sampledBatcher(dev)
        else:
            self.getTrainingBatch, self.numTrainingBatches = self.train.sampledBatcher(batchSize, "label-uniform", cuda, device)
            self.getTestBatch, self.numTestBatches = self.test.sampledBatcher(batchSize, "permutation", cuda, device)

    def close(self) -> None:
        for nb in self._native:
            nb.close()
        self._native = []
