"""Primary / backup coordinator pair: failure detection, takeover and hand-back.

Kept from the reference (src/server.py:183-264):
* the primary pings the backup once per second with ``CheckIfPrimaryUp(req)``; ``req`` is "1" on
  the first ping after a (re)start - "I am recovering" - and "0" afterwards;
* the backup serves ``SendModel`` (stores the replicated global model under ``Backup/``) and
  ``CheckIfPrimaryUp``; a watchdog declares the primary dead when a whole window (10 s by default)
  passes without a ping, and the backup then takes over the round loop;
* a ping with ``req == "1"`` that reaches a backup acting as primary makes it hand the role back;
* sending SIGUSR1 to the backup process toggles its role by hand (operator interface).

Fixed on purpose:
  B1/B2  demotion is a state transition of a small state machine run by ONE thread; the acting
         coordinator is stopped cooperatively and fenced by a term bump, nothing runs inside a
         signal handler and there is no ``Thread.terminate``;
  B6     promotion reloads the replicated model and continues at the next round, not round 0;
  terms  every promotion / recovery increments the term; workers refuse older terms, so two
         coordinators can never drive the same worker at once.
"""
from __future__ import annotations

import os
import queue
import signal
import threading
import time
from dataclasses import dataclass, replace
from typing import Callable, Optional

import grpc

from .. import proto
from ..utils import checkpoint as ckpt
from ..utils.metrics import JsonlSink, log
from . import descriptor as desc
from .coordinator import Coordinator, CoordinatorConfig, GLOBAL_MODEL_FILE


# ------------------------------------------------------------------------------------ primary side
class PrimaryPinger:
    """1 Hz ``CheckIfPrimaryUp`` from the primary to the backup (reference src/server.py:188-200)."""

    def __init__(self, target: str, get_fence: Callable[[], tuple], period_s: float = 1.0,
                 timeout_s: float = 2.0):
        self.channel = proto.make_channel(target)
        self.stub = proto.TrainerStub(self.channel)
        self.get_fence = get_fence
        self.period_s, self.timeout_s = period_s, timeout_s
        self.recovering = 1
        self.stop_event = threading.Event()
        self.pings_ok = 0
        self.thread: Optional[threading.Thread] = None

    def ping_once(self):
        """Returns (ok, backup_term, backup_next_round).  Clears ``recovering`` after the first
        attempt whatever its outcome, like the reference."""
        term, rnd = self.get_fence()
        try:
            _, call = self.stub.CheckIfPrimaryUp.with_call(
                proto.PingRequest(req=str(self.recovering)), timeout=self.timeout_s,
                metadata=desc.fence_metadata(term, rnd))
            bterm, bround = desc.read_fence(call.trailing_metadata())
            self.pings_ok += 1
            return True, bterm, bround
        except grpc.RpcError:
            return False, None, None
        finally:
            self.recovering = 0

    def _loop(self):
        while not self.stop_event.wait(self.period_s):
            self.ping_once()

    def start(self):
        self.thread = threading.Thread(target=self._loop, daemon=True, name="fedb200-pinger")
        self.thread.start()

    def stop(self):
        self.stop_event.set()
        if self.thread:
            self.thread.join(timeout=self.period_s + self.timeout_s + 1)
        self.channel.close()


def run_primary(cfg: CoordinatorConfig, ping_period_s: float = 1.0, sink: Optional[JsonlSink] = None,
                on_coordinator: Optional[Callable[[Coordinator], None]] = None) -> Coordinator:
    """Primary coordinator process body: announce to the backup (taking the role back if the backup
    had been promoted), then run the round loop while pinging."""
    term, next_round, state = Coordinator.load_persisted(cfg.state_dir)
    fence = {"term": term + 1, "round": next_round}
    pinger = None
    if cfg.backup_target:
        pinger = PrimaryPinger(cfg.backup_target, lambda: (fence["term"], fence["round"]), ping_period_s)
        ok, bterm, bround = pinger.ping_once()       # recovering=1: the backup demotes if acting
        if ok and bterm is not None and bterm >= fence["term"]:
            fence["term"] = bterm + 1                # newer than anything the backup handed out
            if bround is not None and bround > next_round:
                # the backup committed rounds while we were away: its workers hold that state
                next_round, state = bround, None
    coord = Coordinator(cfg, term=fence["term"], start_round=next_round, global_state=state, sink=sink)
    fence["round"] = next_round
    if pinger:
        pinger.get_fence = lambda: (coord.term, coord.round)
        pinger.start()
    if on_coordinator:
        on_coordinator(coord)
    try:
        coord.run()
    finally:
        if pinger:
            pinger.stop()
        coord.close()
    return coord


# ------------------------------------------------------------------------------------ backup side
@dataclass
class BackupConfig:
    port: str = "8080"                    # reference --backupPort
    watchdog_s: float = 10.0              # reference src/server.py:260
    state_dir: str = "Backup"
    coordinator: CoordinatorConfig = None  # template for the loop run after promotion


class BackupNode(proto.TrainerServicer):
    """Warm standby.  Roles: BACKUP (serving, watching) <-> ACTING (running the round loop)."""

    def __init__(self, cfg: BackupConfig, sink: Optional[JsonlSink] = None):
        self.cfg = cfg
        self.sink = sink or JsonlSink(None)
        self.events: "queue.Queue[str]" = queue.Queue()
        self.role = "BACKUP"
        self.term = 0
        self.next_round = 0                 # first round not yet replicated to us
        self.last_ping = time.monotonic()
        self.ping_seen = True
        self.acting: Optional[Coordinator] = None
        self._acting_thread: Optional[threading.Thread] = None
        self.stop_event = threading.Event()
        self.promotions = self.demotions = 0
        self.detect_latency_s: Optional[float] = None
        os.makedirs(cfg.state_dir, exist_ok=True)
        t, r, _ = Coordinator.load_persisted(cfg.state_dir)
        self.term, self.next_round = t, r
        self.server: Optional[grpc.Server] = None

    # ------------------------------------------------------------------ rpcs
    def SendModel(self, request, context):
        """Replica store (reference src/server.py:236-242) + which round/term it belongs to."""
        term, rnd = desc.read_fence(context.invocation_metadata())
        if desc.is_descriptor(request.model):
            # GPU plane: the replica itself was written into the backup rank's aux arena by K-REPL / K-AGG over
            # NVLink; the RPC only tells us which round it belongs to
            if term is not None:
                self.term = max(self.term, term)
            if rnd is not None:
                self.next_round = rnd + 1
            self._write_meta()
            self.sink.write("replica", round=rnd, term=term, bytes=0, plane="peer")
            return proto.SendModelReply(reply="success")
        raw = ckpt.decode_wire(request.model)
        path = os.path.join(self.cfg.state_dir, GLOBAL_MODEL_FILE)
        tmp = path + ".tmp"
        with open(tmp, "wb") as f:
            f.write(raw)
        os.replace(tmp, path)
        if term is not None:
            self.term = max(self.term, term)
        if rnd is not None:
            self.next_round = rnd + 1
        self._write_meta()
        self.sink.write("replica", round=rnd, term=term, bytes=len(raw))
        return proto.SendModelReply(reply="success")

    def CheckIfPrimaryUp(self, request, context):
        term, rnd = desc.read_fence(context.invocation_metadata())
        self.last_ping = time.monotonic()
        self.ping_seen = True
        # answer with OUR fence so a recovering primary can pick a newer term / the right round
        my_round = self.acting.round if self.acting is not None else self.next_round
        context.set_trailing_metadata(desc.fence_metadata(self.term, my_round, role=self.role.lower()))
        if request.req == "1" and self.role == "ACTING":
            log.info("recovering primary announced itself: handing the role back")
            self._demote()                   # synchronous: the primary may start as soon as we reply
            context.set_trailing_metadata(desc.fence_metadata(self.term, self.next_round, role="backup"))
        return proto.PingResponse(value=1)

    # ------------------------------------------------------------------ role machine
    def _write_meta(self):
        import json
        tmp = os.path.join(self.cfg.state_dir, f"coordinator.json.tmp.{os.getpid()}.{threading.get_ident()}")
        with open(tmp, "w") as f:
            json.dump({"term": self.term, "next_round": self.next_round,
                       "model": self.cfg.coordinator.model if self.cfg.coordinator else None}, f)
        os.replace(tmp, os.path.join(self.cfg.state_dir, "coordinator.json"))

    def _promote(self):
        if self.role in ("ACTING", "PROMOTING"):
            return
        self.role = "PROMOTING"           # "ACTING" only once self.acting exists: observers read the two together
        self.promotions += 1
        self.term += 1
        _, _, state = Coordinator.load_persisted(self.cfg.state_dir)      # B6: reload the replica
        ccfg = replace(self.cfg.coordinator, state_dir=self.cfg.state_dir, backup_target=None)
        self.acting = Coordinator(ccfg, term=self.term, start_round=self.next_round, global_state=state,
                                  sink=self.sink)
        self.role = "ACTING"
        self.sink.write("promoted", term=self.term, start_round=self.next_round)
        log.warning("primary is down: acting as primary, term %d, from round %d", self.term, self.next_round)

        def body(coord=self.acting):
            try:
                coord.run()
            finally:
                self.next_round = coord.round
                self._write_meta()

        self._acting_thread = threading.Thread(target=body, daemon=True, name="fedb200-acting")
        self._acting_thread.start()

    def _demote(self):
        if self.role != "ACTING":
            return
        coord = self.acting
        coord.stop()
        self.next_round = coord.round
        self.term = max(self.term, coord.term)
        self.role = "BACKUP"
        self.demotions += 1
        self.acting = None
        self.last_ping = time.monotonic()
        self.ping_seen = True
        self._write_meta()
        self.sink.write("demoted", term=self.term, next_round=self.next_round)
        # the loop thread finishes its in-flight RPCs on its own; workers fence it by term anyway
        threading.Thread(target=coord.close, daemon=True).start()

    def _watchdog(self):
        """Reference semantics (src/server.py:254-264): clear the flag, sleep one window, if no
        ping arrived meanwhile the primary is gone.  Detection latency is in (window, 2*window]."""
        while not self.stop_event.is_set():
            if self.role == "BACKUP":
                if self.ping_seen:
                    self.ping_seen = False
                else:
                    self.detect_latency_s = time.monotonic() - self.last_ping
                    self.events.put("promote")
            if self.stop_event.wait(self.cfg.watchdog_s):
                return

    def _machine(self):
        while not self.stop_event.is_set():
            try:
                ev = self.events.get(timeout=0.2)
            except queue.Empty:
                continue
            if ev == "promote":
                self._promote()
            elif ev == "demote":
                self._demote()
            elif ev == "toggle":             # SIGUSR1
                self._demote() if self.role == "ACTING" else self._promote()

    # ------------------------------------------------------------------ lifecycle
    def start(self, install_signal: bool = False):
        self.server = proto.make_server(10, compress=False)   # uncompressed, as the reference
        proto.add_TrainerServicer_to_server(self, self.server)
        if self.server.add_insecure_port(f"[::]:{self.cfg.port}") == 0:
            raise OSError(f"cannot bind backup port {self.cfg.port}")
        self.server.start()
        self.last_ping = time.monotonic()
        for fn, name in ((self._watchdog, "watchdog"), (self._machine, "role-machine")):
            threading.Thread(target=fn, daemon=True, name=f"fedb200-{name}").start()
        if install_signal:
            signal.signal(signal.SIGUSR1, lambda *_: self.events.put("toggle"))
        log.info("backup coordinator listening on :%s (watchdog %.1fs)", self.cfg.port, self.cfg.watchdog_s)

    def wait(self):
        self.server.wait_for_termination()

    def stop(self):
        self.stop_event.set()
        if self.acting is not None:
            self.acting.stop()
        if self.server is not None:
            self.server.stop(grace=0.5)
