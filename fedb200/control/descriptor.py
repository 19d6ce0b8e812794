"""What travels in the proto's string fields.

Two payload families share ``TrainReply.message`` and ``SendModelRequest.model``:

* legacy  - base64(torch.save({'net','acc','epoch'})), byte-compatible with the reference
            (src/client.py:19-23, src/server.py:66-71).  A zip archive in base64 starts "UEsD".
* descriptor - a short ASCII record ``fedb200/1 key=value;key=value`` used when the weights move
            over the GPU peer plane and gRPC only carries the "go"/"done" edges.

Fencing data (term, round) rides in gRPC call metadata, NOT in the messages, so the messages stay
exactly what a reference peer expects (``PingRequest.req`` is still just "1" or "0").
"""
from __future__ import annotations

from typing import Dict, Iterable, Optional, Tuple

MAGIC = "fedb200/1 "
MD_TERM = "fedb200-term"
MD_ROUND = "fedb200-round"
MD_ROLE = "fedb200-role"
MD_LIVE = "fedb200-live"      # comma separated peer ranks that take part in this round (GPU plane)


def encode(kind: str, **fields) -> str:
    parts = [f"kind={kind}"]
    for k, v in fields.items():
        if v is None:
            continue
        s = repr(v) if isinstance(v, float) else str(v)
        if ";" in s or "=" in s:
            raise ValueError(f"descriptor value for {k!r} may not contain ';' or '='")
        parts.append(f"{k}={s}")
    return MAGIC + ";".join(parts)


def is_descriptor(text) -> bool:
    if isinstance(text, bytes):
        return text.startswith(MAGIC.encode())
    return isinstance(text, str) and text.startswith(MAGIC)


def decode(text) -> Dict[str, str]:
    if isinstance(text, bytes):
        text = text.decode("ascii")
    if not is_descriptor(text):
        raise ValueError("not a fedb200 descriptor")
    out: Dict[str, str] = {}
    for item in text[len(MAGIC):].split(";"):
        if item:
            k, _, v = item.partition("=")
            out[k] = v
    return out


def fence_metadata(term: int, round_id: int, role: str = "primary") -> Tuple[Tuple[str, str], ...]:
    return ((MD_TERM, str(term)), (MD_ROUND, str(round_id)), (MD_ROLE, role))


def read_fence(metadata: Optional[Iterable]) -> Tuple[Optional[int], Optional[int]]:
    """(term, round) from invocation/trailing metadata; (None, None) for a reference peer."""
    term = rnd = None
    for item in metadata or ():
        k, v = item[0], item[1]
        if k == MD_TERM:
            term = int(v)
        elif k == MD_ROUND:
            rnd = int(v)
    return term, rnd


def live_metadata(live) -> Tuple[Tuple[str, str], ...]:
    return ((MD_LIVE, ",".join(str(int(r)) for r in live)),)


def read_live(metadata: Optional[Iterable]):
    for item in metadata or ():
        if item[0] == MD_LIVE:
            return [int(x) for x in item[1].split(",") if x != ""]
    return None
