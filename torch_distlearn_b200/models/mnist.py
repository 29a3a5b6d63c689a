"""MNIST convnet of examples/mnist.lua:53-81.

input 1024 -> Reshape(1,32,32) -> conv 1->16 5x5 -> Tanh -> maxpool 2x2 -> conv 16->16 5x5 -> Tanh
-> maxpool 2x2 -> Reshape(400) -> Linear 400->10 -> logSoftMax.   10,842 fp32 parameters in 6
tensors (416 + 6,416 + 4,010)."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class MnistConvNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 16, 5)      # examples/mnist.lua:57
        self.conv2 = nn.Conv2d(16, 16, 5)     # :61
        self.linear = nn.Linear(400, 10)      # :66

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = x.reshape(-1, 1, 32, 32)          # :53 Reshape(1,32,32) of a 1024-vector
        x = F.max_pool2d(torch.tanh(self.conv1(x)), 2, 2)   # :58-59
        x = F.max_pool2d(torch.tanh(self.conv2(x)), 2, 2)   # :62-63
        x = self.linear(x.reshape(x.shape[0], 400))          # :65-66
        return F.log_softmax(x, dim=1)        # :79


def mnist_convnet(seed: int = 0) -> MnistConvNet:
    """Same initial weights on every node: the reference seeds before building (examples/Model.lua:17)."""
    gen_state = torch.random.get_rng_state()
    torch.manual_seed(seed)
    m = MnistConvNet()
    torch.random.set_rng_state(gen_state)
    return m
