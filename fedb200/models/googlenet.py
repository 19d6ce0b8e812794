"""GoogLeNet (CIFAR).  Parity target: reference src/models/googlenet.py:7-99.

Each Inception branch is a Sequential of conv(bias)->BN->ReLU triples (``b4`` starts with a
3x3/1 max-pool), so keys read ``a3.b2.3.weight`` etc.  Branch outputs are concatenated along
channels by ``F.cat`` (K11).
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import AvgPool2d, BatchNorm2d, Conv2d, Linear, MaxPool2d, ReLU, flatten


def _triples(*specs, head=()):
    """specs = (cin, cout, k) triples -> flat module list conv,bn,relu,conv,bn,relu,..."""
    mods = list(head)
    for cin, cout, k in specs:
        mods += [Conv2d(cin, cout, kernel_size=k, padding=k // 2), BatchNorm2d(cout), ReLU(True)]
    return nn.Sequential(*mods)


def _run(seq, x):
    mods, i = list(seq), 0
    while i < len(mods):
        if isinstance(mods[i], Conv2d):
            x = mods[i + 1](mods[i](x), act="relu")
            i += 3
        else:
            x = mods[i](x)
            i += 1
    return x


class Inception(nn.Module):
    def __init__(self, cin, n1x1, n3x3red, n3x3, n5x5red, n5x5, pool_planes):
        super().__init__()
        self.b1 = _triples((cin, n1x1, 1))
        self.b2 = _triples((cin, n3x3red, 1), (n3x3red, n3x3, 3))
        self.b3 = _triples((cin, n5x5red, 1), (n5x5red, n5x5, 3), (n5x5, n5x5, 3))
        self.b4 = _triples((cin, pool_planes, 1), head=(MaxPool2d(3, stride=1, padding=1),))

    def forward(self, x):
        return F.cat([_run(b, x) for b in (self.b1, self.b2, self.b3, self.b4)], 1)


class GoogLeNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.pre_layers = _triples((3, 192, 3))
        self.a3 = Inception(192, 64, 96, 128, 16, 32, 32)
        self.b3 = Inception(256, 128, 128, 192, 32, 96, 64)
        self.maxpool = MaxPool2d(3, stride=2, padding=1)
        self.a4 = Inception(480, 192, 96, 208, 16, 48, 64)
        self.b4 = Inception(512, 160, 112, 224, 24, 64, 64)
        self.c4 = Inception(512, 128, 128, 256, 24, 64, 64)
        self.d4 = Inception(512, 112, 144, 288, 32, 64, 64)
        self.e4 = Inception(528, 256, 160, 320, 32, 128, 128)
        self.a5 = Inception(832, 256, 160, 320, 32, 128, 128)
        self.b5 = Inception(832, 384, 192, 384, 48, 128, 128)
        self.avgpool = AvgPool2d(8, stride=1)
        self.linear = Linear(1024, 10)

    def forward(self, x):
        x = _run(self.pre_layers, x)
        x = self.b3(self.a3(x))
        x = self.maxpool(x)
        for cell in (self.a4, self.b4, self.c4, self.d4, self.e4):
            x = cell(x)
        x = self.maxpool(x)
        x = self.b5(self.a5(x))
        x = self.avgpool(x)
        return self.linear(flatten(x))
