"""NativeBatcher — batches assembled by C++ worker threads (csrc/loader.cpp) into a ring of pinned host buffers.

The reference gets its batches from the ``dataset`` package (Lua over C): ``Dataset(path, {partition, partitions})`` and
``sampledBatcher{samplerKind, batchSize, inputDims, processor}`` -> ``getBatch()`` / ``numBatches()``
(examples/mnist.lua:26-40, examples/cifar10.lua:41-71, examples/Data.lua:22-61).  Here the dataset is a tensor in host
memory (uint8 like the CIFAR-10 binaries, or float32), the sampler is a counter-based generator (batch b is a pure
function of the seed and b, whatever the number of threads), uint8 -> float32 conversion and per-channel normalisation
happen in the worker threads, and the result lands in page-locked memory so that the H2D copy is one asynchronous DMA
(``TrainStep.prefetch`` hides it behind the running step).

    nb = NativeBatcher(images_u8, labels, batch_size=128, sampler="label-uniform", mean=..., std=..., partition=i, partitions=N)
    getBatch, numBatches = nb.sampledBatcher()          # the reference's pair
    batch = getBatch()                                  # Batch(input, target, batchSize): views of a pinned slot
    ...
    nb.close()

A slot is recycled when the consumer has asked for ``hold`` more batches; a slot whose tensors were handed to an
asynchronous copy is protected by the CUDA event passed to ``copied(event)`` (``to_device`` does both).
"""
from __future__ import annotations

import ctypes as C
from collections import deque
from typing import Optional, Sequence, Tuple

import torch

from .. import _lib
from .data import Batch

_KINDS = {"permutation": 0, "label-uniform": 1}


class NativeBatcher:
    def __init__(self, data: torch.Tensor, labels: torch.Tensor, batch_size: int, sampler: str = "permutation",
                 num_classes: Optional[int] = None, mean: Optional[Sequence[float]] = None, std: Optional[Sequence[float]] = None,
                 scale: float = 1.0, partition: int = 1, partitions: int = 1, seed: int = 0, threads: int = 2, depth: int = 4,
                 hold: int = 2, pin: Optional[bool] = None, timeout_s: float = 60.0):
        if sampler not in _KINDS:
            raise ValueError(f"unknown sampler {sampler!r} (permutation | label-uniform)")
        if data.dtype not in (torch.uint8, torch.float32) or data.device.type != "cpu":
            raise TypeError("NativeBatcher: data must be a CPU uint8 or float32 tensor")
        if not 1 <= hold < depth:
            raise ValueError("NativeBatcher: need 1 <= hold < depth")
        self._data = data.contiguous()
        self._labels = labels.to(torch.int64).contiguous()
        n_total = self._data.shape[0]
        if self._labels.numel() != n_total:
            raise ValueError("NativeBatcher: one label per sample")
        self.sample_shape = tuple(self._data.shape[1:])
        elems = int(self._data[0].numel())
        channels = len(mean) if mean is not None else (len(std) if std is not None else 1)
        if elems % channels:
            raise ValueError("NativeBatcher: the number of channels must divide the sample size")
        # x = raw * scale; x = (x - mean) / std   ==   raw * (scale / std) + (-mean / std)
        mean_t = [0.0] * channels if mean is None else [float(m) for m in mean]
        std_t = [1.0] * channels if std is None else [float(s) for s in std]
        sc = (C.c_float * channels)(*[scale / s for s in std_t])
        sh = (C.c_float * channels)(*[-m / s for m, s in zip(mean_t, std_t)])
        self.batch_size, self.depth, self.hold, self.timeout_s = int(batch_size), int(depth), int(hold), float(timeout_s)
        self.num_classes = int(num_classes if num_classes is not None else int(self._labels.max()) + 1)
        pin = torch.cuda.is_available() if pin is None else pin
        self._x = [torch.empty((batch_size,) + self.sample_shape, dtype=torch.float32, pin_memory=pin) for _ in range(depth)]
        self._y = [torch.empty(batch_size, dtype=torch.int64, pin_memory=pin) for _ in range(depth)]
        xs = (C.c_void_p * depth)(*[t.data_ptr() for t in self._x])
        ys = (C.c_void_p * depth)(*[t.data_ptr() for t in self._y])
        self._lib = _lib.host()
        h = C.c_void_p()
        rc = self._lib.dll_create(self._data.data_ptr(), int(self._data.dtype == torch.uint8), self._labels.data_ptr(), n_total, elems,
                                  channels, C.cast(sc, C.c_void_p), C.cast(sh, C.c_void_p), self.num_classes, partition, partitions,
                                  _KINDS[sampler], batch_size, seed, threads, depth, xs, ys, C.byref(h))
        if rc != 0:
            raise RuntimeError(self._lib.dll_last_error().decode())
        self._h = h
        self._out: deque = deque()          # outstanding batches, oldest first: [slot, event or None]
        self.batches = 0

    # ------------------------------------------------------------------------------------------------
    def __len__(self) -> int:
        return int(self._lib.dll_num_samples(self._h))

    def numBatches(self) -> int:
        """Batches per epoch (the last one of a permutation epoch is short when the batch size does not divide n)."""
        return int(self._lib.dll_num_batches(self._h))

    @property
    def samples_done(self) -> int:
        """Samples the worker threads have assembled so far (they run ahead of the consumer by up to ``depth`` batches)."""
        return int(self._lib.dll_samples_done(self._h))

    def _recycle(self) -> None:
        while len(self._out) > self.hold - 1:
            _slot, ev = self._out.popleft()
            if ev is not None:
                ev.synchronize()            # the asynchronous copy out of the slot must be over before a worker rewrites it
            if self._lib.dll_release(self._h) != 0:
                raise RuntimeError(self._lib.dll_last_error().decode())

    def next(self) -> Tuple[torch.Tensor, torch.Tensor, int]:
        """(input, target, batch number): views of the pinned slot, valid until ``hold`` more batches have been taken."""
        if self._h is None:
            raise RuntimeError("NativeBatcher is closed")
        self._recycle()
        slot, size, no = C.c_int(), C.c_int(), C.c_int64()
        rc = self._lib.dll_next(self._h, C.byref(slot), C.byref(size), C.byref(no), self.timeout_s)
        if rc != 0:
            raise RuntimeError(self._lib.dll_last_error().decode())
        self._out.append([slot.value, None])
        self.batches += 1
        return self._x[slot.value][:size.value], self._y[slot.value][:size.value], int(no.value)

    def copied(self, event) -> None:
        """Attach the CUDA event that marks the end of the asynchronous copy out of the batch returned last."""
        if self._out:
            self._out[-1][1] = event

    def to_device(self, device, non_blocking: bool = True) -> Batch:
        """Next batch on ``device``: one asynchronous copy per tensor out of the pinned slot, guarded by an event."""
        x, y, _ = self.next()
        dev = torch.device(device)
        if dev.type != "cuda":
            return Batch(x.clone(), y.clone(), int(x.shape[0]))
        xd, yd = x.to(dev, non_blocking=non_blocking), y.to(dev, non_blocking=non_blocking)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(dev))
        self.copied(ev)
        return Batch(xd, yd, int(x.shape[0]))

    def sampledBatcher(self, device=None):
        """``(getBatch, numBatches)`` like the reference's batcher.  Without a device the batch is a view of the pinned slot."""
        def getBatch() -> Batch:
            if device is not None:
                return self.to_device(device)
            x, y, _ = self.next()
            return Batch(x, y, int(x.shape[0]))
        return getBatch, self.numBatches

    def close(self) -> None:
        if getattr(self, "_h", None) is not None:
            for _slot, ev in self._out:
                if ev is not None:
                    ev.synchronize()
            self._lib.dll_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass


def read_cifar10_bin(paths: Sequence[str], max_records: int = 10000) -> Tuple[torch.Tensor, torch.Tensor]:
    """CIFAR-10 binary batches (``data_batch_N.bin`` / ``test_batch.bin``: 1 label byte + 3072 pixel bytes per record)
    -> (uint8 [n, 3, 32, 32], int64 [n]).  The reference reads the same data from hard-coded home directories
    (examples/cifar10.lua:41,47)."""
    lib = _lib.host()
    xs, ys = [], []
    for p in paths:
        x = torch.empty(max_records, 3072, dtype=torch.uint8)
        y = torch.empty(max_records, dtype=torch.int64)
        n = int(lib.dll_read_cifar_bin(str(p).encode(), x.data_ptr(), y.data_ptr(), max_records))
        if n < 0:
            raise OSError(lib.dll_last_error().decode() + f": {p}")
        xs.append(x[:n])
        ys.append(y[:n])
    return torch.cat(xs).view(-1, 3, 32, 32), torch.cat(ys)
