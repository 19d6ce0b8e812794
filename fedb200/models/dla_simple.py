"""SimpleDLA.  Parity target: reference src/models/dla_simple.py:16-112.

Binary aggregation trees: ``root(cat(left(x), right(left(x))))`` with a 1x1 conv root.
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import BatchNorm2d, Conv2d, Linear, ReLU, flatten
from .resnet import BasicBlock


class Root(nn.Module):
    def __init__(self, cin, cout, kernel_size=1):
        super().__init__()
        self.conv = Conv2d(cin, cout, kernel_size, stride=1, padding=(kernel_size - 1) // 2, bias=False)
        self.bn = BatchNorm2d(cout)

    def forward(self, xs):
        return self.bn(self.conv(F.cat(xs, 1)), act="relu")


class Tree(nn.Module):
    def __init__(self, block, cin, cout, level=1, stride=1):
        super().__init__()
        self.root = Root(2 * cout, cout)
        if level == 1:
            self.left_tree = block(cin, cout, stride=stride)
            self.right_tree = block(cout, cout, stride=1)
        else:
            self.left_tree = Tree(block, cin, cout, level=level - 1, stride=stride)
            self.right_tree = Tree(block, cout, cout, level=level - 1, stride=1)

    def forward(self, x):
        a = self.left_tree(x)
        b = self.right_tree(a)
        return self.root([a, b])


def stem(cin, cout):
    return nn.Sequential(Conv2d(cin, cout, kernel_size=3, stride=1, padding=1, bias=False),
                         BatchNorm2d(cout), ReLU(True))


def run_stem(seq, x):
    return seq[1](seq[0](x), act="relu")


class SimpleDLA(nn.Module):
    def __init__(self, block=BasicBlock, num_classes=10):
        super().__init__()
        self.base = stem(3, 16)
        self.layer1 = stem(16, 16)
        self.layer2 = stem(16, 32)
        self.layer3 = Tree(block, 32, 64, level=1, stride=1)
        self.layer4 = Tree(block, 64, 128, level=2, stride=2)
        self.layer5 = Tree(block, 128, 256, level=2, stride=2)
        self.layer6 = Tree(block, 256, 512, level=1, stride=2)
        self.linear = Linear(512, num_classes)

    def forward(self, x):
        for s in (self.base, self.layer1, self.layer2):
            x = run_stem(s, x)
        for t in (self.layer3, self.layer4, self.layer5, self.layer6):
            x = t(x)
        x = F.avg_pool2d(x, 4)
        return self.linear(flatten(x))
