"""Dual Path Networks DPN-26 / DPN-92.  Parity target: reference src/models/dpn.py:7-89.

A block's output has a residual part (first ``out_planes`` channels, summed with the
shortcut) and a dense part (the rest, concatenated) - ``F.dual_path_merge``.
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import BatchNorm2d, Conv2d, Identity, Linear, flatten, projection


class Bottleneck(nn.Module):
    def __init__(self, last_planes, in_planes, out_planes, dense_depth, stride, first_layer):
        super().__init__()
        self.out_planes, self.dense_depth = out_planes, dense_depth
        self.conv1 = Conv2d(last_planes, in_planes, kernel_size=1, bias=False)
        self.bn1 = BatchNorm2d(in_planes)
        self.conv2 = Conv2d(in_planes, in_planes, kernel_size=3, stride=stride, padding=1, groups=32, bias=False)
        self.bn2 = BatchNorm2d(in_planes)
        self.conv3 = Conv2d(in_planes, out_planes + dense_depth, kernel_size=1, bias=False)
        self.bn3 = BatchNorm2d(out_planes + dense_depth)
        self.shortcut = projection(last_planes, out_planes + dense_depth, stride) if first_layer else Identity()

    def forward(self, x):
        y = self.bn1(self.conv1(x), act="relu")
        y = self.bn2(self.conv2(y), act="relu")
        y = self.bn3(self.conv3(y))
        return F.dual_path_merge(self.shortcut(x), y, self.out_planes)


class DPN(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.conv1 = Conv2d(3, 64, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn1 = BatchNorm2d(64)
        last, stages = 64, []
        for in_p, out_p, n, dd, stride in zip(cfg["in_planes"], cfg["out_planes"], cfg["num_blocks"],
                                              cfg["dense_depth"], (1, 2, 2, 2)):
            mods = []
            for i, s in enumerate([stride] + [1] * (n - 1)):
                mods.append(Bottleneck(last, in_p, out_p, dd, s, i == 0))
                last = out_p + (i + 2) * dd
            stages.append(nn.Sequential(*mods))
        self.last_planes = last
        self.layer1, self.layer2, self.layer3, self.layer4 = stages
        self.linear = Linear(cfg["out_planes"][3] + (cfg["num_blocks"][3] + 1) * cfg["dense_depth"][3], 10)

    def forward(self, x):
        x = self.bn1(self.conv1(x), act="relu")
        for layer in (self.layer1, self.layer2, self.layer3, self.layer4):
            x = layer(x)
        x = F.avg_pool2d(x, 4)
        return self.linear(flatten(x))


def _cfg(num_blocks):
    return {"in_planes": (96, 192, 384, 768), "out_planes": (256, 512, 1024, 2048),
            "num_blocks": num_blocks, "dense_depth": (16, 32, 24, 128)}


def DPN26():
    return DPN(_cfg((2, 2, 2, 2)))


def DPN92():
    return DPN(_cfg((3, 4, 20, 3)))
