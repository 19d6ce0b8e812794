"""ResNeXt-29 family.  Parity target: reference src/models/resnext.py:11-87.

Three stages; the bottleneck width doubles after every stage.  The grouped 3x3 (K4) has
``cardinality`` groups of ``bottleneck_width`` channels.
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import BatchNorm2d, Conv2d, Identity, Linear, flatten, projection


class Block(nn.Module):
    expansion = 2

    def __init__(self, cin, cardinality=32, bottleneck_width=4, stride=1):
        super().__init__()
        gw = cardinality * bottleneck_width
        cout = self.expansion * gw
        self.conv1 = Conv2d(cin, gw, kernel_size=1, bias=False)
        self.bn1 = BatchNorm2d(gw)
        self.conv2 = Conv2d(gw, gw, kernel_size=3, stride=stride, padding=1, groups=cardinality, bias=False)
        self.bn2 = BatchNorm2d(gw)
        self.conv3 = Conv2d(gw, cout, kernel_size=1, bias=False)
        self.bn3 = BatchNorm2d(cout)
        self.shortcut = projection(cin, cout, stride) if (stride != 1 or cin != cout) else Identity()

    def forward(self, x):
        y = self.bn1(self.conv1(x), act="relu")
        y = self.bn2(self.conv2(y), act="relu")
        y = self.bn3(self.conv3(y))
        return F.add(y, self.shortcut(x), act="relu")


class ResNeXt(nn.Module):
    def __init__(self, num_blocks, cardinality, bottleneck_width, num_classes=10):
        super().__init__()
        self.cardinality = cardinality
        self.conv1 = Conv2d(3, 64, kernel_size=1, bias=False)
        self.bn1 = BatchNorm2d(64)
        cin, width, stages = 64, bottleneck_width, []
        for n, stride in zip(num_blocks, (1, 2, 2)):
            mods = []
            for s in [stride] + [1] * (n - 1):
                mods.append(Block(cin, cardinality, width, s))
                cin = Block.expansion * cardinality * width
            stages.append(nn.Sequential(*mods))
            width *= 2
        self.bottleneck_width = width
        self.layer1, self.layer2, self.layer3 = stages
        self.linear = Linear(cardinality * bottleneck_width * 8, num_classes)

    def forward(self, x):
        x = self.bn1(self.conv1(x), act="relu")
        for layer in (self.layer1, self.layer2, self.layer3):
            x = layer(x)
        x = F.avg_pool2d(x, 8)
        return self.linear(flatten(x))


def ResNeXt29_2x64d():
    return ResNeXt((3, 3, 3), cardinality=2, bottleneck_width=64)


def ResNeXt29_4x64d():
    return ResNeXt((3, 3, 3), cardinality=4, bottleneck_width=64)


def ResNeXt29_8x64d():
    return ResNeXt((3, 3, 3), cardinality=8, bottleneck_width=64)


def ResNeXt29_32x4d():
    return ResNeXt((3, 3, 3), cardinality=32, bottleneck_width=4)
