"""CIFAR-10 convnet of examples/Model.lua:20-66 (identical in examples/cifar10.lua:108-156).

4 x [conv 5x5 pad 2 -> SpatialBatchNormalization(eps 1e-3) -> ReLU -> maxpool 2x2/2] with channels
3->64->128->256->512, Reshape(2048), Linear 2048->10, LogSoftMax + ClassNLLCriterion.
4,328,970 fp32 parameters (16.5 MiB) in 18 tensors; BatchNorm running statistics are buffers, not
parameters, so — as in the reference — they are never synchronised between nodes."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


def _execution_order(model: nn.Module):
    """Parameter names in the order the forward pass uses them: conv_i, bn_i, ..., linear (FlatParams(order="execution"))."""
    names = []
    for i in range(len(model.convs)):
        names += [f"convs.{i}.{n}" for n, _ in model.convs[i].named_parameters()]
        names += [f"bns.{i}.{n}" for n, _ in model.bns[i].named_parameters()]
    return names + [f"linear.{n}" for n, _ in model.linear.named_parameters()]


class CifarConvNet(nn.Module):
    CHANNELS = (3, 64, 128, 256, 512)

    def __init__(self):
        super().__init__()
        convs, bns = [], []
        for cin, cout in zip(self.CHANNELS[:-1], self.CHANNELS[1:]):
            convs.append(nn.Conv2d(cin, cout, 5, 1, 2))            # examples/Model.lua:20,26,32,38
            bns.append(nn.BatchNorm2d(cout, eps=1e-3))             # :21,27,33,39
        self.convs = nn.ModuleList(convs)
        self.bns = nn.ModuleList(bns)
        self.linear = nn.Linear(512 * 2 * 2, 10)                   # :44-45

    def execution_order(self):
        return _execution_order(self)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        for conv, bn in zip(self.convs, self.bns):
            x = F.max_pool2d(F.relu(bn(conv(x))), 2, 2)            # :59-61
        x = self.linear(x.reshape(x.shape[0], 2048))               # :63
        return F.log_softmax(x, dim=1)                             # :64


class CifarConvNetTC(nn.Module):
    """The same network with every convolution on the tensor cores (ops/conv.py: im2col + the tcgen05
    3xTF32 GEMM of csrc/tc_gemm.cu, fp32-level accuracy) and NHWC activations.  Same parameter count
    (4,328,970); conv weights are stored as GEMM operands [cout, 25*cin], the Linear layer sees the
    features in the reference's (c, h, w) order, so ``load_reference`` maps parameters one to one."""
    CHANNELS = CifarConvNet.CHANNELS

    def __init__(self, terms: int = 3, fused_tail: bool = True):
        super().__init__()
        from ..ops.bn import FusedBNReLUPool
        from ..ops.conv import TCConv5x5
        self.fused_tail = fused_tail
        self.convs = nn.ModuleList([TCConv5x5(cin, cout, terms) for cin, cout in zip(self.CHANNELS[:-1], self.CHANNELS[1:])])
        self.bns = nn.ModuleList([FusedBNReLUPool(c, eps=1e-3) if fused_tail else nn.BatchNorm2d(c, eps=1e-3)
                                  for c in self.CHANNELS[1:]])
        self.linear = nn.Linear(512 * 2 * 2, 10)
        if fused_tail:
            for conv, bn in zip(self.convs, self.bns):
                bn.attach_conv_bias(conv.bias)                     # bias gradient out of the block's own backward pass

    def execution_order(self):
        return _execution_order(self)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = x.permute(0, 2, 3, 1).contiguous()                     # NCHW input batch -> NHWC once
        for conv, bn in zip(self.convs, self.bns):
            if self.fused_tail:
                x = bn(conv(x))                                    # BN + ReLU + 2x2 max-pool in one pair of passes (csrc/bn_ops.cu)
            else:
                y = conv(x).permute(0, 3, 1, 2)                    # NHWC data seen as a channels_last NCHW tensor
                y = F.max_pool2d(F.relu(bn(y)), 2, 2)              # cuDNN / ATen channels_last kernels
                x = y.permute(0, 2, 3, 1)                          # contiguous NHWC again (no copy)
        x = self.linear(x.permute(0, 3, 1, 2).reshape(x.shape[0], 2048))   # (c, h, w) feature order, examples/Model.lua:44
        return F.log_softmax(x, dim=1)

    def features(self, x: torch.Tensor) -> torch.Tensor:
        """NCHW input batch -> NHWC feature map [B, 2, 2, 512] of the last block (fused tail only)."""
        x = x.permute(0, 2, 3, 1).contiguous()
        for conv, bn in zip(self.convs, self.bns):
            x = bn(conv(x))
        return x

    def forward_with_loss(self, x: torch.Tensor, target: torch.Tensor):
        """(log-probabilities, mean NLL loss) with the classifier head fused into two launches (csrc/head_ops.cu);
        falls back to ``nll_loss(self(x), target)`` where the fused head does not apply (CPU, eval, unfused tail)."""
        from ..ops.head import head_nll, head_supported
        if self.fused_tail and self.training and x.is_cuda:
            feat = self.features(x)
            if head_supported(feat, self.linear.weight):
                loss, logp = head_nll(feat, self.linear.weight, self.linear.bias, target)
                return logp, loss
            out = F.log_softmax(self.linear(feat.permute(0, 3, 1, 2).reshape(feat.shape[0], 2048)), dim=1)
            return out, F.nll_loss(out, target)
        out = self(x)
        return out, F.nll_loss(out, target)

    @torch.no_grad()
    def load_reference(self, ref: CifarConvNet) -> "CifarConvNetTC":
        for mine, theirs in zip(self.convs, ref.convs):
            mine.load_conv2d(theirs)
        for mine, theirs in zip(self.bns, ref.bns):
            mine.load_state_dict(theirs.state_dict())
        self.linear.load_state_dict(ref.linear.state_dict())
        return self


def cifar_convnet_tc(seed: int = 0, terms: int = 3, fused_tail: bool = True) -> CifarConvNetTC:
    """Same initial weights as ``cifar_convnet(seed)`` (built from it), convolutions on tcgen05."""
    return CifarConvNetTC(terms, fused_tail).load_reference(cifar_convnet(seed))


def cifar_convnet(seed: int = 0) -> CifarConvNet:
    """``torch.manualSeed(0)`` on every node before construction (examples/Model.lua:17)."""
    gen_state = torch.random.get_rng_state()
    torch.manual_seed(seed)
    m = CifarConvNet()
    torch.random.set_rng_state(gen_state)
    return m
