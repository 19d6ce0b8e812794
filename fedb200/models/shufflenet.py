"""ShuffleNet G2/G3.  Parity target: reference src/models/shufflenet.py:12-100.

Divergence: the reference computes ``mid_planes = out_planes/4`` as a float and
its constructors raise TypeError on any modern torch; integer division is used here, which
is the evident intent, so both constructors work.
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import AvgPool2d, BatchNorm2d, Conv2d, Identity, Linear, flatten


class ShuffleBlock(nn.Module):
    def __init__(self, groups):
        super().__init__()
        self.groups = groups

    def forward(self, x):
        return F.channel_shuffle(x, self.groups)


class Bottleneck(nn.Module):
    def __init__(self, cin, cout, stride, groups):
        super().__init__()
        self.stride = stride
        mid = cout // 4
        g = 1 if cin == 24 else groups
        self.conv1 = Conv2d(cin, mid, kernel_size=1, groups=g, bias=False)
        self.bn1 = BatchNorm2d(mid)
        self.shuffle1 = ShuffleBlock(groups=g)
        self.conv2 = Conv2d(mid, mid, kernel_size=3, stride=stride, padding=1, groups=mid, bias=False)
        self.bn2 = BatchNorm2d(mid)
        self.conv3 = Conv2d(mid, cout, kernel_size=1, groups=groups, bias=False)
        self.bn3 = BatchNorm2d(cout)
        self.shortcut = nn.Sequential(AvgPool2d(3, stride=2, padding=1)) if stride == 2 else Identity()

    def forward(self, x):
        y = self.shuffle1(self.bn1(self.conv1(x), act="relu"))
        y = self.bn2(self.conv2(y), act="relu")
        y = self.bn3(self.conv3(y))
        res = self.shortcut(x)
        if self.stride == 2:
            return F.relu(F.cat([y, res], 1))
        return F.add(y, res, act="relu")


class ShuffleNet(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        groups = cfg["groups"]
        self.conv1 = Conv2d(3, 24, kernel_size=1, bias=False)
        self.bn1 = BatchNorm2d(24)
        cin, stages = 24, []
        for cout, n in zip(cfg["out_planes"], cfg["num_blocks"]):
            mods = []
            for i in range(n):
                cat_planes = cin if i == 0 else 0
                mods.append(Bottleneck(cin, cout - cat_planes, stride=2 if i == 0 else 1, groups=groups))
                cin = cout
            stages.append(nn.Sequential(*mods))
        self.in_planes = cin
        self.layer1, self.layer2, self.layer3 = stages
        self.linear = Linear(cfg["out_planes"][2], 10)

    def forward(self, x):
        x = self.bn1(self.conv1(x), act="relu")
        for layer in (self.layer1, self.layer2, self.layer3):
            x = layer(x)
        x = F.avg_pool2d(x, 4)
        return self.linear(flatten(x))


def ShuffleNetG2():
    return ShuffleNet({"out_planes": (200, 400, 800), "num_blocks": (4, 8, 4), "groups": 2})


def ShuffleNetG3():
    return ShuffleNet({"out_planes": (240, 480, 960), "num_blocks": (4, 8, 4), "groups": 3})
