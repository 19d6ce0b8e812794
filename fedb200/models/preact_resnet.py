"""Pre-activation ResNets.  Parity target: reference src/models/preact_resnet.py:12-110.

BN -> ReLU -> conv ordering; the projection shortcut (a bare 1x1 conv, no BN) exists only on
blocks that change shape - there is no ``shortcut`` attribute otherwise, as in the reference.
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import BatchNorm2d, Conv2d, Linear, flatten, projection
from .resnet import stage


class PreActBlock(nn.Module):
    expansion = 1

    def __init__(self, cin, planes, stride=1):
        super().__init__()
        self.bn1 = BatchNorm2d(cin)
        self.conv1 = Conv2d(cin, planes, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn2 = BatchNorm2d(planes)
        self.conv2 = Conv2d(planes, planes, kernel_size=3, stride=1, padding=1, bias=False)
        if stride != 1 or cin != self.expansion * planes:
            self.shortcut = projection(cin, self.expansion * planes, stride, with_bn=False)

    def forward(self, x):
        a = self.bn1(x, act="relu")
        skip = self.shortcut(a) if hasattr(self, "shortcut") else x
        y = self.conv1(a)
        y = self.conv2(self.bn2(y, act="relu"))
        return F.add(y, skip)


class PreActBottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin, planes, stride=1):
        super().__init__()
        self.bn1 = BatchNorm2d(cin)
        self.conv1 = Conv2d(cin, planes, kernel_size=1, bias=False)
        self.bn2 = BatchNorm2d(planes)
        self.conv2 = Conv2d(planes, planes, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn3 = BatchNorm2d(planes)
        self.conv3 = Conv2d(planes, self.expansion * planes, kernel_size=1, bias=False)
        if stride != 1 or cin != self.expansion * planes:
            self.shortcut = projection(cin, self.expansion * planes, stride, with_bn=False)

    def forward(self, x):
        a = self.bn1(x, act="relu")
        skip = self.shortcut(a) if hasattr(self, "shortcut") else x
        y = self.conv1(a)
        y = self.conv2(self.bn2(y, act="relu"))
        y = self.conv3(self.bn3(y, act="relu"))
        return F.add(y, skip)


class PreActResNet(nn.Module):
    def __init__(self, block, num_blocks, num_classes=10):
        super().__init__()
        self.conv1 = Conv2d(3, 64, kernel_size=3, stride=1, padding=1, bias=False)
        c = 64
        self.layer1, c = stage(block, c, 64, num_blocks[0], 1)
        self.layer2, c = stage(block, c, 128, num_blocks[1], 2)
        self.layer3, c = stage(block, c, 256, num_blocks[2], 2)
        self.layer4, c = stage(block, c, 512, num_blocks[3], 2)
        self.linear = Linear(c, num_classes)

    def forward(self, x):
        x = self.conv1(x)
        for layer in (self.layer1, self.layer2, self.layer3, self.layer4):
            x = layer(x)
        x = F.avg_pool2d(x, 4)
        return self.linear(flatten(x))


def PreActResNet18():
    return PreActResNet(PreActBlock, (2, 2, 2, 2))


def PreActResNet34():
    return PreActResNet(PreActBlock, (3, 4, 6, 3))


def PreActResNet50():
    return PreActResNet(PreActBottleneck, (3, 4, 6, 3))


def PreActResNet101():
    return PreActResNet(PreActBottleneck, (3, 4, 23, 3))


def PreActResNet152():
    return PreActResNet(PreActBottleneck, (3, 8, 36, 3))
