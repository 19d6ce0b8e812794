"""Observability: structured JSONL per round, a TTY progress bar
that needs no ``stty`` (the reference's utils.py:45 dies without a terminal), NVTX ranges and
CUDA-event phase timers."""
from __future__ import annotations

import contextlib
import json
import logging
import shutil
import sys
import threading
import time
from typing import Any, Dict, Optional

log = logging.getLogger("fedb200")


def setup_logging(level=logging.INFO) -> None:
    if not logging.getLogger().handlers:
        logging.basicConfig(level=level, format="%(asctime)s %(name)s %(levelname)s %(message)s")


class JsonlSink:
    """Append-only JSON-lines file; one record per event; thread safe."""

    def __init__(self, path: Optional[str]):
        self.path = path
        self._lock = threading.Lock()
        self.records = []           # also kept in memory for tests / bench

    def write(self, event: str, **fields: Any) -> Dict[str, Any]:
        rec = {"ts": time.time(), "event": event, **fields}
        with self._lock:
            self.records.append(rec)
            if self.path:
                with open(self.path, "a") as f:
                    f.write(json.dumps(rec, default=str) + "\n")
        return rec


def format_time(seconds: float) -> str:
    """Two most significant units, e.g. '1m12s', '3s250ms', '0ms' (cf. reference utils.py:94-124)."""
    ms = int(round(seconds * 1000))
    units = (("D", 86400000), ("h", 3600000), ("m", 60000), ("s", 1000), ("ms", 1))
    out = []
    for name, size in units:
        q, ms = divmod(ms, size)
        if q and len(out) < 2:
            out.append(f"{q}{name}")
    return "".join(out) or "0ms"


class ProgressBar:
    """``[=====>.....] step/total | Step: 12ms | Tot: 3s | msg``; silent when stdout is not a TTY
    unless ``force``."""

    def __init__(self, total: int, width: int = 40, stream=None, force: bool = False):
        self.total, self.width = max(total, 1), width
        self.stream = stream or sys.stdout
        self.enabled = force or (hasattr(self.stream, "isatty") and self.stream.isatty())
        self.cols = shutil.get_terminal_size((80, 24)).columns
        self.t0 = self.last = time.time()

    def update(self, current: int, msg: str = "") -> str:
        now = time.time()
        step, self.last = now - self.last, now
        done = int(self.width * (current + 1) / self.total)
        bar = "=" * max(done - 1, 0) + ">" + "." * (self.width - done)
        line = (f" [{bar}] {current + 1}/{self.total} | Step: {format_time(step)}"
                f" | Tot: {format_time(now - self.t0)}" + (f" | {msg}" if msg else ""))
        if self.enabled:
            end = "\n" if current + 1 >= self.total else "\r"
            self.stream.write(line[: self.cols - 1].ljust(self.cols - 1) + end)
            self.stream.flush()
        return line


@contextlib.contextmanager
def nvtx_range(name: str):
    """NVTX range when CUDA is up, no-op otherwise."""
    import torch
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


class DeviceTimer:
    """CUDA-event phase timer: ``with t.phase('agg'): ...`` then ``t.ms()`` after a sync."""

    def __init__(self):
        import torch
        self._torch = torch
        self._events = {}

    @contextlib.contextmanager
    def phase(self, name: str, stream=None):
        torch = self._torch
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        try:
            yield
        finally:
            b.record(stream)
            self._events.setdefault(name, []).append((a, b))

    def ms(self, reset: bool = False) -> Dict[str, float]:
        self._torch.cuda.synchronize()
        out = {k: sum(a.elapsed_time(b) for a, b in v) for k, v in self._events.items()}
        if reset:
            self._events = {}
        return out


@contextlib.contextmanager
def device_phase(timer: Optional["DeviceTimer"], name: str, stream=None):
    """``with device_phase(timer, "agg"):`` - an NVTX range plus a CUDA-event pair on ``stream`` when a timer is given
    (profiling on), nothing at all otherwise.  The phases of a federated round (train / agg / repl / eval) are wrapped in
    it by ``GpuFederation`` and ``FedGroup``."""
    if timer is None:
        yield
        return
    with nvtx_range(name), timer.phase(name, stream):
        yield
