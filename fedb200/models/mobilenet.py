"""MobileNet (v1) - the reference's default model (src/main.py:69, src/server.py:158).

state_dict parity target: reference src/models/mobilenet.py:11-52 (164 keys, 27 BN counters).
13 depthwise-separable blocks; on CUDA the depthwise 3x3 is a stencil kernel (K3) and the
pointwise 1x1 is a tcgen05 GEMM over NHWC activations (K1).
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import BatchNorm2d, Conv2d, Linear, flatten

# out_planes or (out_planes, stride)
PLAN = (64, (128, 2), 128, (256, 2), 256, (512, 2), 512, 512, 512, 512, 512, (1024, 2), 1024)


class Block(nn.Module):
    """depthwise 3x3 -> BN -> ReLU -> pointwise 1x1 -> BN -> ReLU"""

    def __init__(self, cin, cout, stride=1):
        super().__init__()
        self.conv1 = Conv2d(cin, cin, kernel_size=3, stride=stride, padding=1, groups=cin, bias=False)
        self.bn1 = BatchNorm2d(cin)
        self.conv2 = Conv2d(cin, cout, kernel_size=1, bias=False)
        self.bn2 = BatchNorm2d(cout)

    def forward(self, x):
        x = self.bn1(self.conv1(x), act="relu")
        return self.bn2(self.conv2(x), act="relu")


class MobileNet(nn.Module):
    cfg = list(PLAN)

    def __init__(self, num_classes=10):
        super().__init__()
        self.conv1 = Conv2d(3, 32, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn1 = BatchNorm2d(32)
        blocks, cin = [], 32
        for item in PLAN:
            cout, stride = (item, 1) if isinstance(item, int) else item
            blocks.append(Block(cin, cout, stride))
            cin = cout
        self.layers = nn.Sequential(*blocks)
        self.linear = Linear(1024, num_classes)

    def forward(self, x):
        x = self.bn1(self.conv1(x), act="relu")
        x = self.layers(x)
        x = F.avg_pool2d(x, 2)
        return self.linear(flatten(x))
