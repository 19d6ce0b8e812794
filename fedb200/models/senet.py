"""SENet-18.  Parity target: reference src/models/senet.py:10-113.

The squeeze-excite gate (global average -> 1x1 conv+ReLU -> 1x1 conv+sigmoid -> channel
scale) runs as ``F.squeeze_excite``; ``SENet18`` uses the pre-activation block, the
post-activation ``BasicBlock`` is kept because the reference exports it.
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import BatchNorm2d, Conv2d, Identity, Linear, flatten, projection


def _se_gate(block, y):
    w = F.global_avg_pool(y)
    w = F.relu(block.fc1(w))
    w = F.sigmoid(block.fc2(w))
    return F.channel_scale(y, w)


class BasicBlock(nn.Module):
    def __init__(self, cin, planes, stride=1):
        super().__init__()
        self.conv1 = Conv2d(cin, planes, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn1 = BatchNorm2d(planes)
        self.conv2 = Conv2d(planes, planes, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn2 = BatchNorm2d(planes)
        self.shortcut = projection(cin, planes, stride) if (stride != 1 or cin != planes) else Identity()
        self.fc1 = Conv2d(planes, planes // 16, kernel_size=1)
        self.fc2 = Conv2d(planes // 16, planes, kernel_size=1)

    def forward(self, x):
        y = self.bn1(self.conv1(x), act="relu")
        y = self.bn2(self.conv2(y))
        y = _se_gate(self, y)
        return F.add(y, self.shortcut(x), act="relu")


class PreActBlock(nn.Module):
    def __init__(self, cin, planes, stride=1):
        super().__init__()
        self.bn1 = BatchNorm2d(cin)
        self.conv1 = Conv2d(cin, planes, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn2 = BatchNorm2d(planes)
        self.conv2 = Conv2d(planes, planes, kernel_size=3, stride=1, padding=1, bias=False)
        if stride != 1 or cin != planes:
            self.shortcut = projection(cin, planes, stride, with_bn=False)
        self.fc1 = Conv2d(planes, planes // 16, kernel_size=1)
        self.fc2 = Conv2d(planes // 16, planes, kernel_size=1)

    def forward(self, x):
        a = self.bn1(x, act="relu")
        skip = self.shortcut(a) if hasattr(self, "shortcut") else x
        y = self.conv1(a)
        y = self.conv2(self.bn2(y, act="relu"))
        y = _se_gate(self, y)
        return F.add(y, skip)


class SENet(nn.Module):
    def __init__(self, block, num_blocks, num_classes=10):
        super().__init__()
        self.conv1 = Conv2d(3, 64, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn1 = BatchNorm2d(64)
        cin, stages = 64, []
        for planes, n, stride in zip((64, 128, 256, 512), num_blocks, (1, 2, 2, 2)):
            mods = []
            for s in [stride] + [1] * (n - 1):
                mods.append(block(cin, planes, s))
                cin = planes
            stages.append(nn.Sequential(*mods))
        self.layer1, self.layer2, self.layer3, self.layer4 = stages
        self.linear = Linear(512, num_classes)

    def forward(self, x):
        x = self.bn1(self.conv1(x), act="relu")
        for layer in (self.layer1, self.layer2, self.layer3, self.layer4):
            x = layer(x)
        x = F.avg_pool2d(x, 4)
        return self.linear(flatten(x))


def SENet18():
    return SENet(PreActBlock, (2, 2, 2, 2))
