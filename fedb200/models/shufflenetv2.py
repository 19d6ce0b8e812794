"""ShuffleNetV2 x0.5/x1/x1.5/x2.  Parity target: reference src/models/shufflenetv2.py:10-152.

BasicBlock: split channels in half, transform the second half, concatenate, shuffle(2).
DownBlock: two stride-2 branches, concatenate, shuffle(2).
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import BatchNorm2d, Conv2d, Linear, flatten

configs = {
    0.5: {"out_channels": (48, 96, 192, 1024), "num_blocks": (3, 7, 3)},
    1: {"out_channels": (116, 232, 464, 1024), "num_blocks": (3, 7, 3)},
    1.5: {"out_channels": (176, 352, 704, 1024), "num_blocks": (3, 7, 3)},
    2: {"out_channels": (224, 488, 976, 2048), "num_blocks": (3, 7, 3)},
}


class ShuffleBlock(nn.Module):
    def __init__(self, groups=2):
        super().__init__()
        self.groups = groups

    def forward(self, x):
        return F.channel_shuffle(x, self.groups)


class SplitBlock(nn.Module):
    def __init__(self, ratio):
        super().__init__()
        self.ratio = ratio

    def forward(self, x):
        c = int(x.size(1) * self.ratio)
        return F.channel_slice(x, 0, c), F.channel_slice(x, c, x.size(1))


class BasicBlock(nn.Module):
    def __init__(self, channels, split_ratio=0.5):
        super().__init__()
        self.split = SplitBlock(split_ratio)
        c = int(channels * split_ratio)
        self.conv1 = Conv2d(c, c, kernel_size=1, bias=False)
        self.bn1 = BatchNorm2d(c)
        self.conv2 = Conv2d(c, c, kernel_size=3, stride=1, padding=1, groups=c, bias=False)
        self.bn2 = BatchNorm2d(c)
        self.conv3 = Conv2d(c, c, kernel_size=1, bias=False)
        self.bn3 = BatchNorm2d(c)
        self.shuffle = ShuffleBlock()

    def forward(self, x):
        keep, work = self.split(x)
        y = self.bn1(self.conv1(work), act="relu")
        y = self.bn2(self.conv2(y))
        y = self.bn3(self.conv3(y), act="relu")
        return self.shuffle(F.cat([keep, y], 1))


class DownBlock(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        mid = cout // 2
        self.conv1 = Conv2d(cin, cin, kernel_size=3, stride=2, padding=1, groups=cin, bias=False)
        self.bn1 = BatchNorm2d(cin)
        self.conv2 = Conv2d(cin, mid, kernel_size=1, bias=False)
        self.bn2 = BatchNorm2d(mid)
        self.conv3 = Conv2d(cin, mid, kernel_size=1, bias=False)
        self.bn3 = BatchNorm2d(mid)
        self.conv4 = Conv2d(mid, mid, kernel_size=3, stride=2, padding=1, groups=mid, bias=False)
        self.bn4 = BatchNorm2d(mid)
        self.conv5 = Conv2d(mid, mid, kernel_size=1, bias=False)
        self.bn5 = BatchNorm2d(mid)
        self.shuffle = ShuffleBlock()

    def forward(self, x):
        left = self.bn1(self.conv1(x))
        left = self.bn2(self.conv2(left), act="relu")
        right = self.bn3(self.conv3(x), act="relu")
        right = self.bn4(self.conv4(right))
        right = self.bn5(self.conv5(right), act="relu")
        return self.shuffle(F.cat([left, right], 1))


class ShuffleNetV2(nn.Module):
    def __init__(self, net_size):
        super().__init__()
        oc, nb = configs[net_size]["out_channels"], configs[net_size]["num_blocks"]
        self.conv1 = Conv2d(3, 24, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn1 = BatchNorm2d(24)
        cin, stages = 24, []
        for cout, n in zip(oc[:3], nb):
            stages.append(nn.Sequential(DownBlock(cin, cout), *[BasicBlock(cout) for _ in range(n)]))
            cin = cout
        self.in_channels = cin
        self.layer1, self.layer2, self.layer3 = stages
        self.conv2 = Conv2d(oc[2], oc[3], kernel_size=1, bias=False)
        self.bn2 = BatchNorm2d(oc[3])
        self.linear = Linear(oc[3], 10)

    def forward(self, x):
        x = self.bn1(self.conv1(x), act="relu")
        for layer in (self.layer1, self.layer2, self.layer3):
            x = layer(x)
        x = self.bn2(self.conv2(x), act="relu")
        x = F.avg_pool2d(x, 4)
        return self.linear(flatten(x))
