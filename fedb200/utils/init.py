"""Parameter initialisation helper of the reference's utils (src/utils.py:29-42), in working form.

The reference's ``init_params`` calls initialisers that were removed from torch.nn.init (``kaiming_normal``, ``constant``,
``normal`` without the trailing underscore) and tests ``if m.bias`` on a tensor, so it cannot run; nothing calls it.  Same
rule here with the current API: Conv2d ~ Kaiming normal over fan_out, BatchNorm2d weight 1 / bias 0, Linear ~ N(0, 1e-3),
every bias 0.  Works on fedb200's layers (they subclass torch.nn) and, because parameters are views into the flat arena,
on a model that has already been adopted by an ``Arena``.
"""
from __future__ import annotations

import torch
from torch import nn


@torch.no_grad()
def init_params(net: nn.Module) -> nn.Module:
    for m in net.modules():
        if isinstance(m, nn.Conv2d):
            nn.init.kaiming_normal_(m.weight, mode="fan_out")
            if m.bias is not None:
                m.bias.zero_()
        elif isinstance(m, nn.BatchNorm2d):
            if m.weight is not None:
                m.weight.fill_(1.0)
                m.bias.zero_()
        elif isinstance(m, nn.Linear):
            m.weight.normal_(0.0, 1e-3)
            if m.bias is not None:
                m.bias.zero_()
    return net
