"""VGG-11/13/16/19 (CIFAR).  Parity target: reference src/models/vgg.py:6-39.

``features`` is one flat Sequential of conv(bias)/BN/ReLU triples and 2x2 max-pools followed by
the reference's AvgPool2d(1,1) no-op, so the key names are ``features.<idx>.*``.  The forward
walks that Sequential and fuses every conv->BN->ReLU triple into one op chain.
"""
import torch.nn as nn

from ..ops.layers import AvgPool2d, BatchNorm2d, Conv2d, Linear, MaxPool2d, ReLU, flatten

cfg = {
    "VGG11": (64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"),
    "VGG13": (64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"),
    "VGG16": (64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"),
    "VGG19": (64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M",
              512, 512, 512, 512, "M"),
}


class VGG(nn.Module):
    def __init__(self, vgg_name):
        super().__init__()
        mods, cin = [], 3
        for item in cfg[vgg_name]:
            if item == "M":
                mods.append(MaxPool2d(kernel_size=2, stride=2))
            else:
                mods += [Conv2d(cin, item, kernel_size=3, padding=1), BatchNorm2d(item), ReLU(inplace=True)]
                cin = item
        mods.append(AvgPool2d(kernel_size=1, stride=1))
        self.features = nn.Sequential(*mods)
        self.classifier = Linear(512, 10)

    def forward(self, x):
        mods, i = list(self.features), 0
        while i < len(mods):
            if isinstance(mods[i], Conv2d):
                x = mods[i + 1](mods[i](x), act="relu")
                i += 3
            else:
                x = mods[i](x)
                i += 1
        return self.classifier(flatten(x))
