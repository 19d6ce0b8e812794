"""PNASNet-A / PNASNet-B.  Parity target: reference src/models/pnasnet.py:10-116.

Cells combine depthwise-style separable convolutions with 7x7 / 5x5 / 3x3 windows
(``groups=in_planes``; out_planes may be a multiple of in_planes on the stride-2 cells) with
a 3x3 max-pool branch.
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import BatchNorm2d, Conv2d, Linear, flatten


class SepConv(nn.Module):
    def __init__(self, cin, cout, kernel_size, stride):
        super().__init__()
        self.conv1 = Conv2d(cin, cout, kernel_size, stride, padding=(kernel_size - 1) // 2,
                            bias=False, groups=cin)
        self.bn1 = BatchNorm2d(cout)

    def forward(self, x):
        return self.bn1(self.conv1(x))


class _Cell(nn.Module):
    def _pool_branch(self, x):
        y = F.max_pool2d(x, kernel_size=3, stride=self.stride, padding=1)
        return self.bn1(self.conv1(y)) if self.stride == 2 else y


class CellA(_Cell):
    def __init__(self, cin, cout, stride=1):
        super().__init__()
        self.stride = stride
        self.sep_conv1 = SepConv(cin, cout, kernel_size=7, stride=stride)
        if stride == 2:
            self.conv1 = Conv2d(cin, cout, kernel_size=1, bias=False)
            self.bn1 = BatchNorm2d(cout)

    def forward(self, x):
        return F.add(self.sep_conv1(x), self._pool_branch(x), act="relu")


class CellB(_Cell):
    def __init__(self, cin, cout, stride=1):
        super().__init__()
        self.stride = stride
        self.sep_conv1 = SepConv(cin, cout, kernel_size=7, stride=stride)
        self.sep_conv2 = SepConv(cin, cout, kernel_size=3, stride=stride)
        self.sep_conv3 = SepConv(cin, cout, kernel_size=5, stride=stride)
        if stride == 2:
            self.conv1 = Conv2d(cin, cout, kernel_size=1, bias=False)
            self.bn1 = BatchNorm2d(cout)
        self.conv2 = Conv2d(2 * cout, cout, kernel_size=1, bias=False)
        self.bn2 = BatchNorm2d(cout)

    def forward(self, x):
        left = F.add(self.sep_conv1(x), self.sep_conv2(x), act="relu")
        right = F.add(self._pool_branch(x), self.sep_conv3(x), act="relu")
        return self.bn2(self.conv2(F.cat([left, right], 1)), act="relu")


class PNASNet(nn.Module):
    def __init__(self, cell_type, num_cells, num_planes):
        super().__init__()
        self.cell_type = cell_type
        self.conv1 = Conv2d(3, num_planes, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn1 = BatchNorm2d(num_planes)
        p = num_planes

        def run(cin, planes):
            return nn.Sequential(*[cell_type(cin if i == 0 else planes, planes, stride=1) for i in range(6)])

        self.layer1 = run(p, p)
        self.layer2 = cell_type(p, 2 * p, stride=2)
        self.layer3 = run(2 * p, 2 * p)
        self.layer4 = cell_type(2 * p, 4 * p, stride=2)
        self.layer5 = run(4 * p, 4 * p)
        self.in_planes = 4 * p
        self.linear = Linear(4 * p, 10)

    def forward(self, x):
        x = self.bn1(self.conv1(x), act="relu")
        for layer in (self.layer1, self.layer2, self.layer3, self.layer4, self.layer5):
            x = layer(x)
        x = F.avg_pool2d(x, 8)
        return self.linear(flatten(x))


def PNASNetA():
    return PNASNet(CellA, num_cells=6, num_planes=44)


def PNASNetB():
    return PNASNet(CellB, num_cells=6, num_planes=32)
