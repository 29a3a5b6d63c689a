"""``optim.ConfusionMatrix`` equivalent (examples/mnist.lua:44,119-125; examples/cifar10.lua:194-205):
``add`` / ``batchAdd`` / ``zero`` / ``.mat`` / ``.totalValid``; the matrix is an int64 tensor so it can
be summed across nodes with the small-message integer all-reduce (``tree.allReduce(cm.mat, add)``)."""
from __future__ import annotations

from typing import Optional, Sequence

import torch


class ConfusionMatrix:
    def __init__(self, classes: Sequence[str], device: Optional[torch.device] = None, comm=None):
        self.classes = list(classes)
        self.nclasses = len(self.classes)
        self.comm = comm
        dev = device or torch.device("cpu")
        if comm is not None:
            # live in the symmetric heap: the cross-node reduce then needs no staging copy
            # (an even element count keeps the int64 buffer a whole number of 16-byte vectors)
            self._buf = comm.alloc((self.nclasses * self.nclasses + 1) // 2 * 2, torch.int64, "confusion")
            self.mat = self._buf.tensor[:self.nclasses * self.nclasses].view(self.nclasses, self.nclasses)
        else:
            self._buf = None
            self.mat = torch.zeros(self.nclasses, self.nclasses, dtype=torch.int64, device=dev)

    def zero(self) -> None:
        self.mat.zero_()

    def add(self, prediction: torch.Tensor, target) -> None:
        """One sample: ``prediction`` is a score vector (or a class index), ``target`` a class index."""
        p = int(prediction.argmax()) if isinstance(prediction, torch.Tensor) and prediction.numel() > 1 else int(prediction)
        self.mat[int(target), p] += 1

    def batchAdd(self, predictions: torch.Tensor, targets: torch.Tensor) -> None:
        """A batch of score rows [B, C] and targets [B]; on CUDA this is one kernel launch."""
        if predictions.is_cuda and self.mat.is_cuda and predictions.dtype == torch.float32 and predictions.is_contiguous():
            from .. import _lib
            lib = _lib.cuda()
            t = targets.to(torch.int64).contiguous()
            with torch.cuda.device(predictions.device):
                _lib.check(lib.dl_confusion_add(predictions.data_ptr(), t.data_ptr(), predictions.shape[0], self.nclasses,
                                                self.mat.data_ptr(), torch.cuda.current_stream().cuda_stream))
            return
        pred = predictions.argmax(dim=1).to("cpu")
        tgt = targets.to("cpu").long()
        idx = tgt * self.nclasses + pred
        self.mat.view(-1).index_add_(0, idx.to(self.mat.device), torch.ones_like(idx, device=self.mat.device))

    def allReduce(self) -> None:
        """Sum over nodes (examples/cifar10.lua:203)."""
        if self.comm is None or self.comm.world == 1:
            return
        self.comm.allreduce(self._buf, self._buf.numel, active=1, normalize=False, variant="oneshot", channel=1, read=False)

    @property
    def totalValid(self) -> float:
        tot = int(self.mat.sum())
        return float(self.mat.diag().sum()) / tot if tot else 0.0

    def __str__(self) -> str:
        m = self.mat.to("cpu")
        lines = ["ConfusionMatrix:"]
        for i, c in enumerate(self.classes):
            row = m[i]
            tot = int(row.sum())
            acc = 100.0 * int(row[i]) / tot if tot else float("nan")
            lines.append("  [" + " ".join(f"{int(v):6d}" for v in row) + f"]  {acc:6.2f}%  [class: {c}]")
        lines.append(f" + global correct: {100.0 * self.totalValid:.3f}%")
        return "\n".join(lines)

    batch_add, all_reduce = batchAdd, allReduce
