"""Flat parameter arena: the live state of a model as two contiguous segments.

The reference moves weights as ``torch.save`` files through disk, base64 and protobuf
(src/client.py:19-23, src/server.py:55-58,155-179).  Here a model's whole ``state_dict`` lives
in ONE fp32 buffer (parameters first, then float buffers such as BN running statistics) plus
ONE int64 buffer (the ``num_batches_tracked`` counters).  Every ``nn.Parameter`` / buffer of the
module is re-pointed to a view of those buffers, so

* FedAvg (K-AGG), broadcast (K-BCAST), replication (K-REPL), the codec (K-CODEC) and SGD (K10)
  are single kernels over flat memory - no per-tensor launches, no pickling;
* on a GPU box the buffers can be VMM allocations mapped into every peer process
  (``fedb200.parallel.mem``), so "uploading" a model is publishing an offset, not sending bytes;
* the manifest (key -> segment, offset, shape, dtype) keeps the on-disk checkpoint layout
  interchangeable with the reference (``fedb200.utils.checkpoint``).

Offsets are aligned to ``ALIGN`` elements (128 B) so every tensor can be read with 16-byte
vector loads and the padding stays zero (harmless under averaging).
"""
from __future__ import annotations

import math
from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, Iterable, List, Optional, Tuple

import torch

ALIGN = 32          # fp32 elements -> 128 bytes
TOTAL_ALIGN = 1024  # whole-segment padding: a multiple of every kernel's vector*unroll width


@dataclass(frozen=True)
class Entry:
    key: str
    segment: str            # "param" | "fbuf" (both inside f32) | "i64"
    offset: int             # element offset inside its buffer (f32 or i64)
    numel: int
    shape: Tuple[int, ...]
    dtype: torch.dtype


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class Manifest:
    """Deterministic layout computed from ``state_dict`` keys/shapes only - every process that
    builds the same model gets the same manifest, which is what lets peers address each other's
    arenas by offset."""

    def __init__(self, entries: List[Entry], n_param: int, n_f32: int, n_i64: int):
        self.entries = entries
        self.by_key: Dict[str, Entry] = {e.key: e for e in entries}
        self.n_param = n_param      # [0, n_param) of f32 = trainable parameters (SGD range)
        self.n_f32 = n_f32          # whole f32 segment (FedAvg range)
        self.n_i64 = n_i64

    @classmethod
    def from_module(cls, module: torch.nn.Module) -> "Manifest":
        param_keys = {k for k, _ in module.named_parameters()}
        sd = module.state_dict()
        entries: Dict[str, Entry] = {}
        off = 0
        for k, t in sd.items():
            if k in param_keys:
                if t.dtype != torch.float32:
                    raise TypeError(f"{k}: parameters must be fp32, got {t.dtype}")
                entries[k] = Entry(k, "param", off, t.numel(), tuple(t.shape), t.dtype)
                off = _round_up(off + t.numel(), ALIGN)
        n_param = _round_up(off, TOTAL_ALIGN)
        off = n_param
        for k, t in sd.items():
            if k not in param_keys and t.dtype.is_floating_point:
                entries[k] = Entry(k, "fbuf", off, t.numel(), tuple(t.shape), t.dtype)
                off = _round_up(off + t.numel(), ALIGN)
        n_f32 = _round_up(off, TOTAL_ALIGN)
        ioff = 0
        for k, t in sd.items():
            if k not in param_keys and not t.dtype.is_floating_point:
                if t.dtype != torch.int64:
                    raise TypeError(f"{k}: integer buffers must be int64, got {t.dtype}")
                entries[k] = Entry(k, "i64", ioff, t.numel(), tuple(t.shape), t.dtype)
                ioff += t.numel()
        n_i64 = _round_up(max(ioff, 1), 16)
        ordered = [entries[k] for k in sd.keys()]   # checkpoint (state_dict) order
        return cls(ordered, n_param, n_f32, n_i64)

    def nbytes(self) -> int:
        return self.n_f32 * 4 + self.n_i64 * 8

    def fingerprint(self) -> str:
        import hashlib
        h = hashlib.sha1()
        for e in self.entries:
            h.update(f"{e.key}|{e.segment}|{e.offset}|{e.shape}|{e.dtype};".encode())
        return h.hexdigest()[:16]


class Arena:
    """A model's state (and, for trainers, gradient + momentum) in flat buffers."""

    def __init__(self, manifest: Manifest, device="cpu", f32: Optional[torch.Tensor] = None,
                 i64: Optional[torch.Tensor] = None):
        self.manifest = manifest
        self.device = torch.device(device)
        self.f32 = f32 if f32 is not None else torch.zeros(manifest.n_f32, dtype=torch.float32, device=device)
        self.i64 = i64 if i64 is not None else torch.zeros(manifest.n_i64, dtype=torch.int64, device=device)
        assert self.f32.numel() >= manifest.n_f32 and self.i64.numel() >= manifest.n_i64
        self.grad: Optional[torch.Tensor] = None       # [n_param], trainers only
        self.momentum: Optional[torch.Tensor] = None   # [n_param]
        self.sgd_steps = 0                             # 0 => first step seeds momentum with g

    # ------------------------------------------------------------------ views
    def view(self, key: str) -> torch.Tensor:
        e = self.manifest.by_key[key]
        buf = self.i64 if e.segment == "i64" else self.f32
        return buf[e.offset:e.offset + e.numel].view(e.shape)

    def grad_view(self, key: str) -> torch.Tensor:
        e = self.manifest.by_key[key]
        assert e.segment == "param" and self.grad is not None
        return self.grad[e.offset:e.offset + e.numel].view(e.shape)

    # ------------------------------------------------------------------ module binding
    @classmethod
    def adopt(cls, module: torch.nn.Module, device=None, with_optimizer_state: bool = False,
              f32: Optional[torch.Tensor] = None, i64: Optional[torch.Tensor] = None) -> "Arena":
        """Moves ``module``'s parameters and buffers into a fresh arena (values preserved) and
        re-points them at arena views.  After this call ``module.state_dict()`` reads the arena
        and any kernel that writes the arena updates the module in place."""
        if device is None:
            device = next(module.parameters()).device
        man = Manifest.from_module(module)
        arena = cls(man, device, f32=f32, i64=i64)
        sd = module.state_dict()
        with torch.no_grad():
            for e in man.entries:
                arena.view(e.key).copy_(sd[e.key])
        if with_optimizer_state:
            arena.grad = torch.zeros(man.n_param, dtype=torch.float32, device=device)
            arena.momentum = torch.zeros(man.n_param, dtype=torch.float32, device=device)
        arena.bind(module)
        return arena

    def bind(self, module: torch.nn.Module) -> None:
        params = dict(module.named_parameters())
        for e in self.manifest.entries:
            v = self.view(e.key)
            if e.key in params:
                p = params[e.key]
                p.data = v
                if self.grad is not None:
                    p.grad = self.grad_view(e.key)
            else:
                mod, _, name = e.key.rpartition(".")
                owner = module.get_submodule(mod) if mod else module
                owner._buffers[name] = v

    # ------------------------------------------------------------------ state_dict <-> arena
    def state_dict(self, to_cpu: bool = True) -> "OrderedDict[str, torch.Tensor]":
        out = OrderedDict()
        for e in self.manifest.entries:
            t = self.view(e.key)
            out[e.key] = t.detach().cpu().clone() if to_cpu else t
        return out

    def load_state_dict(self, sd, strict: bool = True) -> None:
        """Copy semantics of ``nn.Module.load_state_dict``: ``copy_`` converts dtypes, so a
        reference checkpoint whose ``num_batches_tracked`` is fp32 (1.5) lands truncated toward
        zero (1) in the int64 segment - SURVEY N3."""
        keys = set(sd.keys())
        mine = set(self.manifest.by_key)
        if strict and keys != mine:
            raise KeyError(f"state_dict mismatch: missing {sorted(mine - keys)[:4]}, "
                           f"unexpected {sorted(keys - mine)[:4]}")
        with torch.no_grad():
            for e in self.manifest.entries:
                if e.key in sd:
                    self.view(e.key).copy_(sd[e.key])

    def copy_from(self, other: "Arena") -> None:
        self.f32[:self.manifest.n_f32].copy_(other.f32[:self.manifest.n_f32])
        self.i64[:self.manifest.n_i64].copy_(other.i64[:self.manifest.n_i64])

    def zero_grad(self) -> None:
        if self.grad is not None:
            self.grad.zero_()


def split_ranges(n: int, parts: int, granule: int = TOTAL_ALIGN) -> List[Tuple[int, int]]:
    """Contiguous, granule-aligned shards of [0, n) - the two-shot K-AGG ownership map."""
    per = _round_up(math.ceil(n / parts), granule)
    return [(min(i * per, n), min((i + 1) * per, n)) for i in range(parts)]


def assign_roles(live, pref_primary: int = 0, pref_backup: int = 1):
    """(holder of the server's global model, holder of its replica) for a live set on the GPU plane.  The preferred ranks
    while they are live, else the lowest live ranks; with a single live rank both roles fall to it.  Pure function of
    its arguments: every rank (and the coordinator, which names the holder in its ``pull`` descriptor) computes the same
    answer from the same live list.  The reference has no such notion - its server process owns the model - but it does
    re-admit ANY worker (src/server.py:78-101), so no worker rank may be special here."""
    lv = sorted(live)
    if not lv:
        raise ValueError("empty live set")
    primary = pref_primary if pref_primary in lv else lv[0]
    rest = [r for r in lv if r != primary]
    backup = (pref_backup if pref_backup in rest else rest[0]) if rest else primary
    return primary, backup
