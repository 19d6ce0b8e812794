"""MobileNetV2 (CIFAR variant).  Parity target: reference src/models/mobilenetv2.py:11-78."""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import BatchNorm2d, Conv2d, Identity, Linear, flatten, projection

# (expansion, out_planes, num_blocks, stride)
PLAN = ((1, 16, 1, 1), (6, 24, 2, 1), (6, 32, 3, 2), (6, 64, 4, 2),
        (6, 96, 3, 1), (6, 160, 3, 2), (6, 320, 1, 1))


class Block(nn.Module):
    """1x1 expand -> depthwise 3x3 -> 1x1 project, residual when stride == 1"""

    def __init__(self, cin, cout, expansion, stride):
        super().__init__()
        self.stride = stride
        mid = expansion * cin
        self.conv1 = Conv2d(cin, mid, kernel_size=1, bias=False)
        self.bn1 = BatchNorm2d(mid)
        self.conv2 = Conv2d(mid, mid, kernel_size=3, stride=stride, padding=1, groups=mid, bias=False)
        self.bn2 = BatchNorm2d(mid)
        self.conv3 = Conv2d(mid, cout, kernel_size=1, bias=False)
        self.bn3 = BatchNorm2d(cout)
        self.shortcut = projection(cin, cout, 1) if (stride == 1 and cin != cout) else Identity()

    def forward(self, x):
        y = self.bn1(self.conv1(x), act="relu")
        y = self.bn2(self.conv2(y), act="relu")
        y = self.bn3(self.conv3(y))
        return F.add(y, self.shortcut(x)) if self.stride == 1 else y


class MobileNetV2(nn.Module):
    cfg = [tuple(p) for p in PLAN]

    def __init__(self, num_classes=10):
        super().__init__()
        self.conv1 = Conv2d(3, 32, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn1 = BatchNorm2d(32)
        blocks, cin = [], 32
        for expansion, cout, n, stride in PLAN:
            for s in [stride] + [1] * (n - 1):
                blocks.append(Block(cin, cout, expansion, s))
                cin = cout
        self.layers = nn.Sequential(*blocks)
        self.conv2 = Conv2d(320, 1280, kernel_size=1, bias=False)
        self.bn2 = BatchNorm2d(1280)
        self.linear = Linear(1280, num_classes)

    def forward(self, x):
        x = self.bn1(self.conv1(x), act="relu")
        x = self.layers(x)
        x = self.bn2(self.conv2(x), act="relu")
        x = F.avg_pool2d(x, 4)
        return self.linear(flatten(x))
