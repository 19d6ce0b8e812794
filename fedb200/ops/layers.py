"""Layer modules of the fedb200 model zoo.

Every layer subclasses the matching ``torch.nn`` module so that parameter names, shapes,
dtypes, registration order and default initialisation are exactly PyTorch's - which is
what makes ``state_dict`` files interchangeable with the reference.
Only ``forward`` differs: CUDA tensors go to the hand-written sm_100a kernels in
``fedb200.ops.functional``; CPU tensors take the stock PyTorch path, which doubles as the
numerical oracle in the tests and as the compute path of CPU workers (BASELINE config 1).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import functional as F


class Conv2d(nn.Conv2d):
    """K1/K3/K4: dense, depthwise or grouped convolution (reference: every models/*.py)."""

    def forward(self, x):
        return F.conv2d(x, self.weight, self.bias, self.stride, self.padding, self.groups)


class BatchNorm2d(nn.BatchNorm2d):
    """K5: training / eval batch normalisation with PyTorch's running-stat semantics."""

    def forward(self, x, act: str = "none"):
        return F.batch_norm(x, self, act)


class Linear(nn.Linear):
    """K8: classifier GEMM (reference: models/mobilenet.py:35 and friends)."""

    def forward(self, x):
        return F.linear(x, self.weight, self.bias)


class ReLU(nn.ReLU):
    def forward(self, x):
        return F.relu(x)


class MaxPool2d(nn.MaxPool2d):
    def forward(self, x):
        return F.max_pool2d(x, self.kernel_size, self.stride, self.padding)


class AvgPool2d(nn.AvgPool2d):
    def forward(self, x):
        return F.avg_pool2d(x, self.kernel_size, self.stride, self.padding)


class Identity(nn.Sequential):
    """Empty ``nn.Sequential`` stand-in (keeps the attribute, adds no state_dict keys)."""

    def forward(self, x):
        return x


def conv_bn(cin, cout, k, stride=1, padding=0, groups=1, bias=False):
    """Returns the (conv, bn) pair most blocks are made of."""
    return (Conv2d(cin, cout, kernel_size=k, stride=stride, padding=padding, groups=groups,
                   bias=bias), BatchNorm2d(cout))


def projection(cin, cout, stride, with_bn=True):
    """1x1 projection shortcut as ``Sequential(conv[, bn])`` - key names ``shortcut.0/1``."""
    mods = [Conv2d(cin, cout, kernel_size=1, stride=stride, bias=False)]
    if with_bn:
        mods.append(BatchNorm2d(cout))
    return nn.Sequential(*mods)


class ConvBNAct(nn.Sequential):
    """``Sequential(conv, bn, relu)`` triple whose forward runs as one fused op chain."""

    def __init__(self, cin, cout, k, padding=0, bias=True):
        super().__init__(Conv2d(cin, cout, kernel_size=k, padding=padding, bias=bias),
                         BatchNorm2d(cout), ReLU(True))

    def forward(self, x):
        return self[1](self[0](x), act="relu")


def flatten(x: torch.Tensor) -> torch.Tensor:
    """[N,C,H,W] -> [N, C*H*W] in NCHW order whatever the memory format."""
    return x.reshape(x.size(0), -1)
