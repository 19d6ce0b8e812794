"""ResNet-18/34/50/101/152 (CIFAR).  Parity target: reference src/models/resnet.py:14-124.

ResNet-18 is BASELINE config 3.  On CUDA the 3x3 convolutions are tcgen05 implicit GEMMs
whose A tiles are shifted TMA loads of the NHWC activation (fedb200/csrc/conv_tc.cu).
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import BatchNorm2d, Conv2d, Identity, Linear, flatten, projection


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, cin, planes, stride=1):
        super().__init__()
        self.conv1 = Conv2d(cin, planes, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn1 = BatchNorm2d(planes)
        self.conv2 = Conv2d(planes, planes, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn2 = BatchNorm2d(planes)
        cout = self.expansion * planes
        self.shortcut = projection(cin, cout, stride) if (stride != 1 or cin != cout) else Identity()

    def forward(self, x):
        y = self.bn1(self.conv1(x), act="relu")
        y = self.bn2(self.conv2(y))
        return F.add(y, self.shortcut(x), act="relu")


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin, planes, stride=1):
        super().__init__()
        cout = self.expansion * planes
        self.conv1 = Conv2d(cin, planes, kernel_size=1, bias=False)
        self.bn1 = BatchNorm2d(planes)
        self.conv2 = Conv2d(planes, planes, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn2 = BatchNorm2d(planes)
        self.conv3 = Conv2d(planes, cout, kernel_size=1, bias=False)
        self.bn3 = BatchNorm2d(cout)
        self.shortcut = projection(cin, cout, stride) if (stride != 1 or cin != cout) else Identity()

    def forward(self, x):
        y = self.bn1(self.conv1(x), act="relu")
        y = self.bn2(self.conv2(y), act="relu")
        y = self.bn3(self.conv3(y))
        return F.add(y, self.shortcut(x), act="relu")


def stage(block, cin, planes, n, stride):
    """``n`` blocks, the first one strided; returns (Sequential, channels out)."""
    mods = []
    for s in [stride] + [1] * (n - 1):
        mods.append(block(cin, planes, s))
        cin = planes * block.expansion
    return nn.Sequential(*mods), cin


class ResNet(nn.Module):
    def __init__(self, block, num_blocks, num_classes=10):
        super().__init__()
        self.conv1 = Conv2d(3, 64, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn1 = BatchNorm2d(64)
        c = 64
        self.layer1, c = stage(block, c, 64, num_blocks[0], 1)
        self.layer2, c = stage(block, c, 128, num_blocks[1], 2)
        self.layer3, c = stage(block, c, 256, num_blocks[2], 2)
        self.layer4, c = stage(block, c, 512, num_blocks[3], 2)
        self.linear = Linear(c, num_classes)

    def forward(self, x):
        x = self.bn1(self.conv1(x), act="relu")
        for layer in (self.layer1, self.layer2, self.layer3, self.layer4):
            x = layer(x)
        x = F.avg_pool2d(x, 4)
        return self.linear(flatten(x))


def ResNet18():
    return ResNet(BasicBlock, (2, 2, 2, 2))


def ResNet34():
    return ResNet(BasicBlock, (3, 4, 6, 3))


def ResNet50():
    return ResNet(Bottleneck, (3, 4, 6, 3))


def ResNet101():
    return ResNet(Bottleneck, (3, 4, 23, 3))


def ResNet152():
    return ResNet(Bottleneck, (3, 8, 36, 3))
