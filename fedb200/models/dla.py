"""DLA (paper version).  Parity target: reference src/models/dla.py:14-110.

A level-L tree aggregates ``L+2`` feature maps at its root: prev_root, the outputs of the
sub-trees level_{L-1}..level_1 (registered in that descending order), left_node, right_node.
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import Linear, flatten
from .dla_simple import Root, run_stem, stem
from .resnet import BasicBlock


class Tree(nn.Module):
    def __init__(self, block, cin, cout, level=1, stride=1):
        super().__init__()
        self.level = level
        if level == 1:
            self.root = Root(2 * cout, cout)
            self.left_node = block(cin, cout, stride=stride)
            self.right_node = block(cout, cout, stride=1)
        else:
            self.root = Root((level + 2) * cout, cout)
            for i in range(level - 1, 0, -1):
                setattr(self, f"level_{i}", Tree(block, cin, cout, level=i, stride=stride))
            self.prev_root = block(cin, cout, stride=stride)
            self.left_node = block(cout, cout, stride=1)
            self.right_node = block(cout, cout, stride=1)

    def forward(self, x):
        feats = [self.prev_root(x)] if self.level > 1 else []
        for i in range(self.level - 1, 0, -1):
            x = getattr(self, f"level_{i}")(x)
            feats.append(x)
        x = self.left_node(x)
        feats.append(x)
        x = self.right_node(x)
        feats.append(x)
        return self.root(feats)


class DLA(nn.Module):
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
