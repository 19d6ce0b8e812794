"""Deterministic fault injection (SURVEY §5.3 - the reference has none).

A fault plan is a list of rules ``action:phase[:round]`` given through ``FEDB200_FAULTS`` (comma
separated) or ``install()``.  ``fire(phase, round)`` is called by the coordinator at named points
of a round (round_start, after_train, after_agg, after_repl, round_end) and by workers
(worker_train).  Actions: ``kill`` = ``os._exit(137)`` (a crash, no cleanup), ``raise`` = raise
``InjectedFault``, ``sleep<secs>`` = stall.  Each rule fires once.
"""
from __future__ import annotations

import os
import threading
import time
from typing import List, Optional

_lock = threading.Lock()
_rules: Optional[List[dict]] = None


class InjectedFault(RuntimeError):
    pass


def _parse(spec: str) -> List[dict]:
    rules = []
    for item in filter(None, (s.strip() for s in spec.split(","))):
        parts = item.split(":")
        if len(parts) < 2:
            raise ValueError(f"bad fault rule {item!r}; want action:phase[:round]")
        rules.append({"action": parts[0], "phase": parts[1],
                      "round": int(parts[2]) if len(parts) > 2 else None, "done": False})
    return rules


def install(spec: str) -> None:
    global _rules
    with _lock:
        _rules = _parse(spec)


def _ensure():
    global _rules
    if _rules is None:
        _rules = _parse(os.environ.get("FEDB200_FAULTS", ""))


def fire(phase: str, round_id: Optional[int] = None) -> None:
    with _lock:
        _ensure()
        todo = [r for r in _rules if not r["done"] and r["phase"] == phase
                and (r["round"] is None or r["round"] == round_id)]
        for r in todo:
            r["done"] = True
    for r in todo:
        act = r["action"]
        if act == "kill":
            os._exit(137)
        elif act == "raise":
            raise InjectedFault(f"{phase}@{round_id}")
        elif act.startswith("sleep"):
            time.sleep(float(act[5:] or 1))
        else:
            raise ValueError(f"unknown fault action {act!r}")
