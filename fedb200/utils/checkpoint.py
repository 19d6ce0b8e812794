"""Checkpoint format and the legacy wire codec.

File layout, unchanged from the reference (src/main.py:160-165, src/server.py:174-179):
``torch.save({'net': OrderedDict[str -> Tensor], 'acc': number, 'epoch': int}, path)`` with plain
module key names in registration order, fp32 tensors plus int64 BN counters.  Files written here
load in the reference and vice versa.  The legacy data plane ships exactly these bytes as base64
text inside a proto ``string`` (src/client.py:19-23); ``encode_wire`` / ``decode_wire`` do that.

Divergences, on purpose: the directory is created on demand and ':' in a worker address is kept
in the file stem only if the filesystem allows it.
"""
from __future__ import annotations

import base64
import io
import os
from collections import OrderedDict
from typing import Any, Dict

import torch


def make_state(net_state: "OrderedDict[str, torch.Tensor]", acc=1, epoch=1) -> Dict[str, Any]:
    return {"net": net_state, "acc": acc, "epoch": epoch}


def save_checkpoint(path: str, net_state, acc=1, epoch=1) -> None:
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    tmp = f"{path}.tmp.{os.getpid()}"
    torch.save(make_state(net_state, acc, epoch), tmp)
    os.replace(tmp, path)   # readers never observe a half-written file


def load_checkpoint(path_or_bytes, map_location="cpu") -> Dict[str, Any]:
    if isinstance(path_or_bytes, (bytes, bytearray, memoryview)):
        path_or_bytes = io.BytesIO(bytes(path_or_bytes))
    ckpt = torch.load(path_or_bytes, map_location=map_location, weights_only=True)
    if not isinstance(ckpt, dict) or "net" not in ckpt:
        raise ValueError("not a {'net','acc','epoch'} checkpoint")
    return ckpt


def dumps(net_state, acc=1, epoch=1) -> bytes:
    buf = io.BytesIO()
    torch.save(make_state(net_state, acc, epoch), buf)
    return buf.getvalue()


def encode_wire(raw: bytes) -> bytes:
    """checkpoint bytes -> base64 ASCII, assignable to a proto3 string field."""
    return base64.b64encode(raw)


def decode_wire(text) -> bytes:
    if isinstance(text, str):
        text = text.encode("ascii")
    return base64.b64decode(text)


def worker_checkpoint_path(address: str, root: str = "./checkpoint") -> str:
    """``./checkpoint/<address|temp>.pth`` (reference: src/main.py:83-85, src/client.py:19)."""
    return os.path.join(root, f"{address or 'temp'}.pth")
