"""RegNetX-200MF / X-400MF / Y-400MF.  Parity target: reference src/models/regnet.py:11-143."""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import BatchNorm2d, Conv2d, Identity, Linear, flatten, projection


class SE(nn.Module):
    def __init__(self, planes, se_planes):
        super().__init__()
        self.se1 = Conv2d(planes, se_planes, kernel_size=1, bias=True)
        self.se2 = Conv2d(se_planes, planes, kernel_size=1, bias=True)

    def forward(self, x):
        g = F.global_avg_pool(x)
        g = F.relu(self.se1(g))
        g = F.sigmoid(self.se2(g))
        return F.channel_scale(x, g)


class Block(nn.Module):
    def __init__(self, w_in, w_out, stride, group_width, bottleneck_ratio, se_ratio):
        super().__init__()
        w_b = int(round(w_out * bottleneck_ratio))
        self.conv1 = Conv2d(w_in, w_b, kernel_size=1, bias=False)
        self.bn1 = BatchNorm2d(w_b)
        self.conv2 = Conv2d(w_b, w_b, kernel_size=3, stride=stride, padding=1,
                            groups=w_b // group_width, bias=False)
        self.bn2 = BatchNorm2d(w_b)
        self.with_se = se_ratio > 0
        if self.with_se:
            self.se = SE(w_b, int(round(w_in * se_ratio)))
        self.conv3 = Conv2d(w_b, w_out, kernel_size=1, bias=False)
        self.bn3 = BatchNorm2d(w_out)
        self.shortcut = projection(w_in, w_out, stride) if (stride != 1 or w_in != w_out) else Identity()

    def forward(self, x):
        y = self.bn1(self.conv1(x), act="relu")
        y = self.bn2(self.conv2(y), act="relu")
        if self.with_se:
            y = self.se(y)
        y = self.bn3(self.conv3(y))
        return F.add(y, self.shortcut(x), act="relu")


class RegNet(nn.Module):
    def __init__(self, cfg, num_classes=10):
        super().__init__()
        self.cfg = cfg
        self.conv1 = Conv2d(3, 64, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn1 = BatchNorm2d(64)
        cin, stages = 64, []
        for depth, width, stride in zip(cfg["depths"], cfg["widths"], cfg["strides"]):
            mods = []
            for i in range(depth):
                mods.append(Block(cin, width, stride if i == 0 else 1, cfg["group_width"],
                                  cfg["bottleneck_ratio"], cfg["se_ratio"]))
                cin = width
            stages.append(nn.Sequential(*mods))
        self.in_planes = cin
        self.layer1, self.layer2, self.layer3, self.layer4 = stages
        self.linear = Linear(cfg["widths"][-1], num_classes)

    def forward(self, x):
        x = self.bn1(self.conv1(x), act="relu")
        for layer in (self.layer1, self.layer2, self.layer3, self.layer4):
            x = layer(x)
        return self.linear(flatten(F.global_avg_pool(x)))


def _cfg(depths, widths, group_width, se_ratio):
    return {"depths": depths, "widths": widths, "strides": (1, 1, 2, 2), "group_width": group_width,
            "bottleneck_ratio": 1, "se_ratio": se_ratio}


def RegNetX_200MF():
    return RegNet(_cfg((1, 1, 4, 7), (24, 56, 152, 368), 8, 0))


def RegNetX_400MF():
    return RegNet(_cfg((1, 2, 7, 12), (32, 64, 160, 384), 16, 0))


def RegNetY_400MF():
    return RegNet(_cfg((1, 2, 7, 12), (32, 64, 160, 384), 16, 0.25))
