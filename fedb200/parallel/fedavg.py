"""FedAvg - the aggregation semantics of the reference, stated once.

``fedavg_state_dicts`` is a literal CPU restatement of reference src/server.py:163-171 and is
the ORACLE for the CUDA aggregation kernel (K-AGG, csrc/fedavg.cu).  ``aggregate_arenas`` is
the product path: flat buffers in, flat buffers out, CUDA kernel when the arenas are on a GPU.

Semantics:
* N1  unweighted mean in fp32, sequential order ``((p0+p1)+p2...)/N``, division last.
* N2  every key is averaged - weights, biases, BN running_mean/var.
* N3  int64 ``num_batches_tracked`` counters are summed in int64 and true-divided (both
      operands converted to fp32) -> fp32; when that lands in a model the copy truncates
      toward zero, so the arena form stores ``trunc(float(sum)/float(N))`` as int64.
* weighted mode (north star): weights ``n_i`` are per-client sample counts.  If they are all
  equal the uniform formula above is used verbatim (bit-identical to N1); otherwise
  ``(((n0*p0)+(n1*p1))+...)/sum(n)`` in fp32 with the products rounded before the adds.
* only LIVE uploads enter the mean (fixes SURVEY B3).
"""
from __future__ import annotations

from collections import OrderedDict
from typing import List, Optional, Sequence

import torch

from .arena import Arena


def _uniform(weights: Optional[Sequence[int]]) -> bool:
    return weights is None or len(set(int(w) for w in weights)) <= 1


def fedavg_state_dicts(sds: List[dict], weights: Optional[Sequence[int]] = None) -> "OrderedDict":
    """Oracle.  Returns what the reference coordinator writes to ``optimizedModel.pth``['net']:
    fp32 tensors for every key, including the (fractional) averaged BN counters."""
    if not sds:
        raise ValueError("fedavg over zero uploads")
    n = len(sds)
    out = OrderedDict()
    if _uniform(weights):
        for k in sds[0]:
            acc = sds[0][k]
            for i in range(1, n):
                acc = acc + sds[i][k]
            out[k] = acc / n
        return out
    tot = float(sum(int(w) for w in weights))
    for k in sds[0]:
        if sds[0][k].dtype.is_floating_point:
            acc = sds[0][k] * float(weights[0])
            for i in range(1, n):
                acc = acc + sds[i][k] * float(weights[i])
            out[k] = acc / tot
        else:  # counters stay an unweighted mean: they count steps, not samples
            acc = sds[0][k]
            for i in range(1, n):
                acc = acc + sds[i][k]
            out[k] = acc / n
    return out


def fedavg_flat_cpu(f32s: List[torch.Tensor], i64s: List[torch.Tensor],
                    weights: Optional[Sequence[int]], out_f32: torch.Tensor, out_i64: torch.Tensor):
    """Flat-arena form of the oracle (CPU tensors)."""
    n = len(f32s)
    if _uniform(weights):
        acc = f32s[0].clone()
        for i in range(1, n):
            acc += f32s[i]
        out_f32.copy_(acc / n)
    else:
        tot = float(sum(int(w) for w in weights))
        acc = f32s[0] * float(weights[0])
        for i in range(1, n):
            acc = acc + f32s[i] * float(weights[i])
        out_f32.copy_(acc / tot)
    s = i64s[0].clone()
    for i in range(1, n):
        s += i64s[i]
    out_i64.copy_((s.to(torch.float32) / float(n)).trunc().to(torch.int64))


def aggregate_arenas(clients: List[Arena], out: Arena, weights: Optional[Sequence[int]] = None,
                     stream=None) -> None:
    """out <- FedAvg(clients).  All arenas share one manifest.  GPU arenas (local or peer-mapped)
    run K-AGG; CPU arenas run the oracle."""
    man = out.manifest
    if out.f32.is_cuda:
        from ..ops import cuda as K
        K.fedavg(out, clients, weights, stream=stream)
        return
    fedavg_flat_cpu([c.f32[:man.n_f32] for c in clients], [c.i64[:man.n_i64] for c in clients],
                    weights, out.f32[:man.n_f32], out.i64[:man.n_i64])
