"""Functional op surface of fedb200 (K1-K13 of SURVEY §2.6).

Dispatch rule, the same for every op: a CUDA tensor runs the hand-written sm_100a kernel
(``fedb200.ops.cuda``; a missing ``libfedb200.so`` raises, it never silently falls back), a
CPU tensor runs the plain PyTorch op, which is the oracle the kernels are tested against.
"""
from __future__ import annotations

import torch
import torch.nn.functional as TF


def _cuda():
    from . import cuda  # deferred: importing it loads libfedb200.so
    return cuda


def _pair(v):
    return (v, v) if isinstance(v, int) else tuple(v)


# ----------------------------------------------------------------------------- conv / linear
def conv2d(x, weight, bias=None, stride=1, padding=0, groups=1):
    if x.is_cuda:
        return _cuda().conv2d(x, weight, bias, _pair(stride), _pair(padding), groups)
    return TF.conv2d(x, weight, bias, stride, padding, 1, groups)


def linear(x, weight, bias=None):
    if x.is_cuda:
        return _cuda().linear(x, weight, bias)
    return TF.linear(x, weight, bias)


# ----------------------------------------------------------------------------- normalisation
def batch_norm(x, bn, act: str = "none"):
    """BatchNorm2d forward with optional fused activation (``none|relu|swish``).

    Train mode updates ``running_mean/var`` with momentum 0.1 (unbiased variance) and
    increments ``num_batches_tracked`` exactly like ``torch.nn.BatchNorm2d`` - FedAvg
    averages those buffers too.
    """
    if x.is_cuda:
        return _cuda().batch_norm(x, bn, act)
    y = TF.batch_norm(x, bn.running_mean, bn.running_var, bn.weight, bn.bias,
                      bn.training, bn.momentum, bn.eps)
    if bn.training:
        bn.num_batches_tracked += 1
    return activation(y, act)


def activation(x, act: str):
    if act == "none":
        return x
    if act == "relu":
        return relu(x)
    if act == "swish":
        return swish(x)
    if act == "sigmoid":
        return sigmoid(x)
    raise ValueError(act)


# ----------------------------------------------------------------------------- elementwise
def relu(x):
    return _cuda().relu(x) if x.is_cuda else TF.relu(x)


def sigmoid(x):
    return _cuda().sigmoid(x) if x.is_cuda else torch.sigmoid(x)


def swish(x):
    """x * sigmoid(x) (reference: models/efficientnet.py:12-13)."""
    return _cuda().swish(x) if x.is_cuda else x * torch.sigmoid(x)


def add(a, b, act: str = "none"):
    """Residual add with optional fused ReLU."""
    if a.is_cuda:
        return _cuda().add(a, b, act)
    return activation(a + b, act)


def channel_scale(x, s):
    """SE excitation: x[N,C,H,W] * s[N,C,1,1] (reference: models/senet.py:73)."""
    if x.is_cuda:
        return _cuda().channel_scale(x, s)
    return x * s


def drop_connect(x, drop_ratio: float, training: bool):
    """Per-sample stochastic depth (reference: models/efficientnet.py:16-22)."""
    if not training or drop_ratio <= 0:
        return x
    keep = 1.0 - drop_ratio
    if x.is_cuda:
        return _cuda().drop_connect(x, keep)
    mask = torch.empty([x.shape[0], 1, 1, 1], dtype=x.dtype, device=x.device).bernoulli_(keep)
    return x / keep * mask


def dropout(x, p: float, training: bool):
    if not training or p <= 0:
        return x
    if x.is_cuda:
        return _cuda().dropout(x, p)
    return TF.dropout(x, p=p, training=True)


# ----------------------------------------------------------------------------- pooling
def max_pool2d(x, kernel_size, stride=None, padding=0):
    stride = kernel_size if stride is None else stride
    if x.is_cuda:
        return _cuda().pool2d(x, "max", _pair(kernel_size), _pair(stride), _pair(padding))
    return TF.max_pool2d(x, kernel_size, stride, padding)


def avg_pool2d(x, kernel_size, stride=None, padding=0):
    stride = kernel_size if stride is None else stride
    if _pair(kernel_size) == (1, 1) and _pair(stride) == (1, 1):
        return x  # AvgPool2d(1,1) no-op (reference: models/vgg.py:37)
    if x.is_cuda:
        return _cuda().pool2d(x, "avg", _pair(kernel_size), _pair(stride), _pair(padding))
    return TF.avg_pool2d(x, kernel_size, stride, padding)


def global_avg_pool(x):
    """adaptive_avg_pool2d(x, 1)."""
    if x.is_cuda:
        return _cuda().pool2d(x, "avg", tuple(x.shape[2:]), tuple(x.shape[2:]), (0, 0))
    return TF.adaptive_avg_pool2d(x, 1)


# ----------------------------------------------------------------------------- layout moves
def cat(xs, dim=1):
    if xs[0].is_cuda:
        return _cuda().cat_channels(list(xs))
    return torch.cat(list(xs), dim)


def channel_slice(x, lo, hi):
    if x.is_cuda:
        return _cuda().channel_slice(x, lo, hi)
    return x[:, lo:hi]


def channel_shuffle(x, groups):
    """[N,C,H,W] -> [N,g,C/g,H,W] -> transpose(1,2) -> [N,C,H,W]
    (reference: models/shufflenet.py:15-19)."""
    if x.is_cuda:
        return _cuda().channel_shuffle(x, groups)
    n, c, h, w = x.shape
    return x.view(n, groups, c // groups, h, w).permute(0, 2, 1, 3, 4).reshape(n, c, h, w)


# ----------------------------------------------------------------------------- loss
def cross_entropy(logits, target):
    """Mean softmax cross-entropy (reference: src/main.py:98,149)."""
    if logits.is_cuda:
        return _cuda().cross_entropy(logits, target)
    return TF.cross_entropy(logits, target)


# ----------------------------------------------------------------------------- fused chains
def conv_relu_pool(x, weight, bias):
    """valid conv (stride 1, no padding) -> ReLU -> 2x2 max-pool, LeNet's feature stage
    (reference: models/lenet.py:15-18).  One kernel on CUDA."""
    if x.is_cuda:
        return _cuda().conv_relu_pool(x, weight, bias)
    return TF.max_pool2d(TF.relu(TF.conv2d(x, weight, bias)), 2)


def linear_act(x, weight, bias, act: str = "none"):
    if x.is_cuda:
        return _cuda().linear(x, weight, bias, act)
    return activation(TF.linear(x, weight, bias), act)


def dual_path_merge(x, y, d):
    """DPN merge: relu(cat(x[:, :d] + y[:, :d], x[:, d:], y[:, d:]))
    (reference: models/dpn.py:32-35)."""
    if x.is_cuda:
        return _cuda().dual_path_merge(x, y, d)
    return TF.relu(torch.cat([x[:, :d] + y[:, :d], x[:, d:], y[:, d:]], 1))
