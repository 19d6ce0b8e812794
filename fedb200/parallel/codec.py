"""CPU restatement of K-CODEC (csrc/codec.cu), the opt-in lossy meaning of ``-c`` on the NVLink plane.

The reference's ``-c Y`` is lossless gzip of base64 text (src/server.py:103-107, src/client.py:39-43); it has no lossy mode.
This module is the ORACLE of the GPU kernels - same block size, same scale rule, same rounding (e4m3, round to nearest even,
values never exceed the format's maximum by construction), same error feedback - so the kernels can be checked bit for bit, and
CPU-only code can reason about the codec without a GPU.

Per block of 32 consecutive elements:  d = (cur - base) + residual;  s = max|d| / 448 (1 for an all-zero block);
q = e4m3(d / s);  residual <- d - float(q) * s.   Decode:  out = base + float(q) * s.
"""
from __future__ import annotations

from typing import Tuple

import torch

BLOCK = 32
E4M3_MAX = 448.0


def _blocks(t: torch.Tensor) -> torch.Tensor:
    n = t.numel()
    pad = (-n) % BLOCK
    if pad:
        t = torch.cat([t.reshape(-1), t.new_zeros(pad)])
    return t.reshape(-1, BLOCK)


def encode(cur: torch.Tensor, base: torch.Tensor, residual: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Returns (q as uint8 [n], scale fp32 [ceil(n / 32)]) and updates ``residual`` in place."""
    n = cur.numel()
    d = _blocks((cur.reshape(-1).float() - base.reshape(-1).float()) + residual.reshape(-1).float())
    amax = d.abs().amax(dim=1, keepdim=True)
    s = torch.where(amax > 0, amax / E4M3_MAX, torch.ones_like(amax))
    q = (d / s).to(torch.float8_e4m3fn)
    back = q.float() * s
    residual.reshape(-1).copy_((d - back).reshape(-1)[:n])
    return q.view(torch.uint8).reshape(-1)[:n].clone(), s.reshape(-1).clone()


def decode(base: torch.Tensor, q: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    n = base.numel()
    qf = _blocks(q.reshape(-1)).view(torch.float8_e4m3fn).float()
    return base.reshape(-1).float() + (qf * scale.reshape(-1, 1)).reshape(-1)[:n]


def fedavg_decoded(base: torch.Tensor, qs, scales) -> torch.Tensor:
    """What K-AGG computes from compressed uploads: base + ((d_0 + d_1) + ...) / N with the deltas decoded on the fly."""
    n = base.numel()
    acc = None
    for q, s in zip(qs, scales):
        d = (_blocks(q.reshape(-1)).view(torch.float8_e4m3fn).float() * s.reshape(-1, 1)).reshape(-1)[:n]
        acc = d if acc is None else acc + d
    return base.reshape(-1).float() + acc / float(len(qs))
