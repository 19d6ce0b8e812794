"""DenseNet-121/169/201/161 and the small ``densenet_cifar``.
Parity target: reference src/models/densenet.py:9-99.  Every layer concatenates its
``growth_rate`` new channels IN FRONT of its input (``cat([out, x])``), order matters."""
import math

import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import BatchNorm2d, Conv2d, Linear, flatten


class Bottleneck(nn.Module):
    def __init__(self, cin, growth_rate):
        super().__init__()
        self.bn1 = BatchNorm2d(cin)
        self.conv1 = Conv2d(cin, 4 * growth_rate, kernel_size=1, bias=False)
        self.bn2 = BatchNorm2d(4 * growth_rate)
        self.conv2 = Conv2d(4 * growth_rate, growth_rate, kernel_size=3, padding=1, bias=False)

    def forward(self, x):
        y = self.conv1(self.bn1(x, act="relu"))
        y = self.conv2(self.bn2(y, act="relu"))
        return F.cat([y, x], 1)


class Transition(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.bn = BatchNorm2d(cin)
        self.conv = Conv2d(cin, cout, kernel_size=1, bias=False)

    def forward(self, x):
        return F.avg_pool2d(self.conv(self.bn(x, act="relu")), 2)


class DenseNet(nn.Module):
    def __init__(self, block, nblocks, growth_rate=12, reduction=0.5, num_classes=10):
        super().__init__()
        self.growth_rate = growth_rate
        planes = 2 * growth_rate
        self.conv1 = Conv2d(3, planes, kernel_size=3, padding=1, bias=False)
        for i, n in enumerate(nblocks, start=1):
            layers = []
            for _ in range(n):
                layers.append(block(planes, growth_rate))
                planes += growth_rate
            setattr(self, f"dense{i}", nn.Sequential(*layers))
            if i < len(nblocks):
                out = int(math.floor(planes * reduction))
                setattr(self, f"trans{i}", Transition(planes, out))
                planes = out
        self.bn = BatchNorm2d(planes)
        self.linear = Linear(planes, num_classes)

    def forward(self, x):
        x = self.conv1(x)
        x = self.trans1(self.dense1(x))
        x = self.trans2(self.dense2(x))
        x = self.trans3(self.dense3(x))
        x = self.dense4(x)
        x = F.avg_pool2d(self.bn(x, act="relu"), 4)
        return self.linear(flatten(x))


def DenseNet121():
    return DenseNet(Bottleneck, (6, 12, 24, 16), growth_rate=32)


def DenseNet169():
    return DenseNet(Bottleneck, (6, 12, 32, 32), growth_rate=32)


def DenseNet201():
    return DenseNet(Bottleneck, (6, 12, 48, 32), growth_rate=32)


def DenseNet161():
    return DenseNet(Bottleneck, (6, 12, 36, 24), growth_rate=48)


def densenet_cifar():
    return DenseNet(Bottleneck, (6, 12, 24, 16), growth_rate=12)
