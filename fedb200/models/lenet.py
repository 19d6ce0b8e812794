"""LeNet - the zoo's "small CNN" and the flagship of bench.py.

state_dict parity target: reference src/models/lenet.py:5-23 (10 keys, 62,006 params, no BN).
Needs 3x32x32 input: fc1 is 16*5*5 wide.  On CUDA the conv+ReLU+maxpool stages run as one
fused kernel each (fedb200/ops/cuda.py: conv_relu_pool).
"""
import torch.nn as nn

from ..ops import functional as F
from ..ops.layers import Conv2d, Linear, flatten


class LeNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = Conv2d(3, 6, 5)
        self.conv2 = Conv2d(6, 16, 5)
        self.fc1 = Linear(16 * 5 * 5, 120)
        self.fc2 = Linear(120, 84)
        self.fc3 = Linear(84, 10)

    def features(self, x):
        for conv in (self.conv1, self.conv2):
            x = F.conv_relu_pool(x, conv.weight, conv.bias)
        return flatten(x)

    def forward(self, x):
        h = self.features(x)
        h = F.linear_act(h, self.fc1.weight, self.fc1.bias, "relu")
        h = F.linear_act(h, self.fc2.weight, self.fc2.bias, "relu")
        return self.fc3(h)
