"""EfficientNet-B0 (CIFAR).  Parity target: reference src/models/efficientnet.py:12-165.

MBConv blocks with swish, squeeze-excite (swish inside), depthwise 3x3 / 5x5, per-sample
drop-connect on the residual branch (K12) and dropout before the classifier.
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import BatchNorm2d, Conv2d, Linear, flatten

swish = F.swish
drop_connect = lambda x, drop_ratio: F.drop_connect(x, drop_ratio, True)  # noqa: E731


class SE(nn.Module):
    def __init__(self, channels, se_channels):
        super().__init__()
        self.se1 = Conv2d(channels, se_channels, kernel_size=1, bias=True)
        self.se2 = Conv2d(se_channels, channels, kernel_size=1, bias=True)

    def forward(self, x):
        g = F.global_avg_pool(x)
        g = F.swish(self.se1(g))
        g = F.sigmoid(self.se2(g))
        return F.channel_scale(x, g)


class Block(nn.Module):
    def __init__(self, cin, cout, kernel_size, stride, expand_ratio=1, se_ratio=0., drop_rate=0.):
        super().__init__()
        self.stride, self.drop_rate, self.expand_ratio = stride, drop_rate, expand_ratio
        mid = expand_ratio * cin
        self.conv1 = Conv2d(cin, mid, kernel_size=1, bias=False)
        self.bn1 = BatchNorm2d(mid)
        self.conv2 = Conv2d(mid, mid, kernel_size=kernel_size, stride=stride,
                            padding=(1 if kernel_size == 3 else 2), groups=mid, bias=False)
        self.bn2 = BatchNorm2d(mid)
        self.se = SE(mid, int(cin * se_ratio))
        self.conv3 = Conv2d(mid, cout, kernel_size=1, bias=False)
        self.bn3 = BatchNorm2d(cout)
        self.has_skip = (stride == 1) and (cin == cout)

    def forward(self, x):
        # expand_ratio == 1 blocks own conv1/bn1 parameters but never use them (as upstream)
        y = x if self.expand_ratio == 1 else self.bn1(self.conv1(x), act="swish")
        y = self.bn2(self.conv2(y), act="swish")
        y = self.se(y)
        y = self.bn3(self.conv3(y))
        if self.has_skip:
            y = F.drop_connect(y, self.drop_rate, self.training)
            y = F.add(y, x)
        return y


class EfficientNet(nn.Module):
    def __init__(self, cfg, num_classes=10):
        super().__init__()
        self.cfg = cfg
        self.conv1 = Conv2d(3, 32, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn1 = BatchNorm2d(32)
        blocks, cin, b, total = [], 32, 0, sum(cfg["num_blocks"])
        for exp, cout, n, k, stride in zip(cfg["expansion"], cfg["out_channels"], cfg["num_blocks"],
                                           cfg["kernel_size"], cfg["stride"]):
            for s in [stride] + [1] * (n - 1):
                # b is never advanced upstream, so every block gets drop rate 0 (kept for parity)
                blocks.append(Block(cin, cout, k, s, exp, se_ratio=0.25,
                                    drop_rate=cfg["drop_connect_rate"] * b / total))
                cin = cout
        self.layers = nn.Sequential(*blocks)
        self.linear = Linear(cfg["out_channels"][-1], num_classes)

    def forward(self, x):
        x = self.bn1(self.conv1(x), act="swish")
        x = self.layers(x)
        x = flatten(F.global_avg_pool(x))
        x = F.dropout(x, self.cfg["dropout_rate"], self.training)
        return self.linear(x)


def EfficientNetB0():
    return EfficientNet({
        "num_blocks": (1, 2, 2, 3, 3, 4, 1),
        "expansion": (1, 6, 6, 6, 6, 6, 6),
        "out_channels": (16, 24, 40, 80, 112, 192, 320),
        "kernel_size": (3, 3, 5, 3, 5, 5, 3),
        "stride": (1, 2, 2, 2, 1, 2, 1),
        "dropout_rate": 0.2,
        "drop_connect_rate": 0.2,
    })
