"""Coordinator: the FedAvg round loop, worker membership and replication to the backup
(SURVEY C4, C5, C7, C8, C10).

One round, as in the reference (src/server.py:113-153):
  1. ``StartTrain(rank, world)`` on every LIVE worker in parallel; ``rank`` is the dense index among
     live workers, ``world`` the CONFIGURED worker count (N9);
  2. FedAvg over the uploads (``fedb200.parallel.fedavg``);
  3. push the new global model to the backup coordinator;
  4. ``SendModel`` to every live worker in parallel (each evaluates inside the RPC, N8).

Fixed on purpose:
  B3  only uploads of workers that answered THIS round are averaged;
  B5  a failed backup push marks the backup down, it never touches the worker table;
  B6  state (global model, round, term) is persisted and reloaded on restart / promotion;
  B10 workers, rounds, model are configuration; B11 every RPC has a deadline;
  B12 ``run`` returns after the last round and every helper thread is a stoppable daemon.
Fencing: every RPC carries (term, round) metadata; a coordinator that sees FAILED_PRECONDITION
from a worker has been deposed and stops.
"""
from __future__ import annotations

import json
import os
import threading
import time
from collections import OrderedDict
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

import grpc
import torch

from .. import models, proto
from ..parallel.fedavg import fedavg_state_dicts
from ..utils import checkpoint as ckpt
from ..utils.metrics import JsonlSink, log
from . import descriptor as desc
from . import faults

GLOBAL_MODEL_FILE = "optimizedModel.pth"   # reference src/server.py:34
STATE_FILE = "coordinator.json"


@dataclass
class CoordinatorConfig:
    workers: Sequence[str] = ("localhost:50051", "localhost:50052")   # reference src/server.py:281-282
    rounds: int = 20                                                   # reference src/server.py:120
    model: str = "MobileNet"                                           # reference src/server.py:158
    compress: bool = False                                             # -c Y, worker channels only
    state_dir: str = "Primary"                                         # Primary/ or Backup/
    backup_target: Optional[str] = None                                # host:port of the backup
    replicate: str = "async"            # sync: block the round like the reference | async: overlap
    weighted: bool = False              # sample-count weighting (north star); False == N1
    keep_uploads: bool = True           # also write test_<rank>.pth like the reference
    train_timeout_s: float = 3600.0
    send_timeout_s: float = 3600.0
    heartbeat_timeout_s: float = 2.0
    membership_period_s: float = 1.0
    metrics_path: Optional[str] = None
    plane: str = "legacy"               # legacy: weights in the RPC payloads | peer: weights over NVLink, RPCs
                                        # carry descriptors (workers must run with --plane peer on one box)
    checkpoint_every: int = 0           # peer plane: every k rounds ask the primary's rank to dump
                                        # optimizedModel.pth (0 = never; the global model lives in GPU memory)


class Membership:
    """Worker liveness table.  Single writer per transition, all access under one lock (the
    reference mutates bare dicts from N+2 threads)."""

    def __init__(self, workers: Sequence[str], compress: bool):
        self._lock = threading.RLock()
        self.order = list(workers)
        self.compress = compress
        self.live: Dict[str, bool] = {w: True for w in self.order}
        self.channels: Dict[str, grpc.Channel] = {w: proto.make_channel(w, compress) for w in self.order}
        self.events: List[tuple] = []

    def live_workers(self) -> List[str]:
        with self._lock:
            return [w for w in self.order if self.live[w]]

    def dead_workers(self) -> List[str]:
        with self._lock:
            return [w for w in self.order if not self.live[w]]

    def stub(self, w: str) -> proto.TrainerStub:
        with self._lock:
            return proto.TrainerStub(self.channels[w])

    def mark_dead(self, w: str, why: str = "") -> None:
        with self._lock:
            if self.live.get(w):
                self.live[w] = False
                self.events.append((time.time(), w, "down", why))
                log.warning("worker %s marked down (%s)", w, why)

    def readmit(self, w: str, channel: grpc.Channel) -> None:
        with self._lock:
            old = self.channels.get(w)
            self.channels[w] = channel
            self.live[w] = True
            self.events.append((time.time(), w, "up", ""))
        if old is not None and old is not channel:
            old.close()
        log.info("worker %s re-admitted", w)

    def snapshot(self) -> Dict[str, bool]:
        with self._lock:
            return dict(self.live)

    def close(self):
        with self._lock:
            for ch in self.channels.values():
                ch.close()


class Coordinator:
    def __init__(self, cfg: CoordinatorConfig, term: int = 1, start_round: int = 0,
                 global_state: Optional[OrderedDict] = None, sink: Optional[JsonlSink] = None):
        self.cfg = cfg
        self.term = term
        self.round = start_round
        self.sink = sink or JsonlSink(cfg.metrics_path)
        self.members = Membership(cfg.workers, cfg.compress)
        self.global_state: Optional[OrderedDict] = global_state
        self.stop_event = threading.Event()
        self.deposed = False
        self.backup_up = cfg.backup_target is not None
        self._backup_channel = (proto.make_channel(cfg.backup_target, compress=False)
                                if cfg.backup_target else None)   # uncompressed, as the reference
        self._repl_thread: Optional[threading.Thread] = None
        self._member_thread: Optional[threading.Thread] = None
        self.round_times: List[float] = []
        # peer plane: a round died inside the device-flag protocol -> put the survivors on a fresh epoch first.
        # A coordinator that takes over mid-run (promotion, restart) cannot know what is in flight: it resets too.
        self._need_reset = cfg.plane == "peer" and start_round > 0
        os.makedirs(cfg.state_dir, exist_ok=True)

    # ------------------------------------------------------------------ persistence (C10, B6)
    def _path(self, name: str) -> str:
        return os.path.join(self.cfg.state_dir, name)

    def persist_model(self) -> None:
        if self.global_state is not None:
            ckpt.save_checkpoint(self._path(GLOBAL_MODEL_FILE), self.global_state, acc=1, epoch=1)

    def persist_meta(self) -> None:
        """``next_round`` = first round that has NOT been committed yet."""
        tmp = self._path(STATE_FILE) + f".tmp.{os.getpid()}.{threading.get_ident()}"
        with open(tmp, "w") as f:
            json.dump({"term": self.term, "next_round": self.round, "model": self.cfg.model}, f)
        os.replace(tmp, self._path(STATE_FILE))

    def persist(self) -> None:
        self.persist_model()
        self.persist_meta()

    @staticmethod
    def load_persisted(state_dir: str):
        """(term, next round to run, global state or None) from ``state_dir``."""
        term, rnd, state = 0, 0, None
        try:
            with open(os.path.join(state_dir, STATE_FILE)) as f:
                meta = json.load(f)
            term, rnd = int(meta["term"]), int(meta["next_round"])
        except (OSError, ValueError, KeyError):
            pass
        p = os.path.join(state_dir, GLOBAL_MODEL_FILE)
        if os.path.exists(p):
            try:
                state = ckpt.load_checkpoint(p)["net"]
            except Exception as e:   # torn file from a crash mid-write of a foreign writer
                log.warning("ignoring unreadable %s: %s", p, e)
        return term, rnd, state

    # ------------------------------------------------------------------ membership (C7)
    def _membership_loop(self):
        while not self.stop_event.wait(self.cfg.membership_period_s):
            for w in self.members.dead_workers():
                ch = proto.make_channel(w, self.cfg.compress)
                try:
                    rep = proto.TrainerStub(ch).HeartBeat(proto.Request(), timeout=self.cfg.heartbeat_timeout_s)
                except grpc.RpcError:
                    ch.close()
                    continue
                if rep.status == 1:
                    if self.cfg.plane == "peer":
                        # the newcomer is a new process with new GPU memory: survivors map it again, then it
                        # pulls the global model out of the primary's peer memory
                        r = list(self.cfg.workers).index(w)
                        remap = desc.encode("remap", rank=r)
                        survivors = self.members.live_workers()
                        for other in survivors:
                            self._send_model_to(other, remap)
                        self.members.readmit(w, ch)
                        if self.round > 0 and survivors:
                            # the lowest live rank holds the global model (fedb200/parallel/group.py: _assign_roles)
                            from ..parallel.arena import assign_roles
                            holder, _ = assign_roles([list(self.cfg.workers).index(o) for o in survivors])
                            self._send_model_to(w, desc.encode("pull", src=holder))
                        self._need_reset = True
                        continue
                    self.members.readmit(w, ch)
                    if self.global_state is not None:       # hand the newcomer the current model
                        self._send_model_to(w)
            self.sink.write("membership", live=self.members.snapshot(), term=self.term)

    def start_membership(self):
        if self._member_thread is None:
            self._member_thread = threading.Thread(target=self._membership_loop, daemon=True,
                                                   name="fedb200-membership")
            self._member_thread.start()

    # ------------------------------------------------------------------ phases
    def _metadata(self):
        return desc.fence_metadata(self.term, self.round)

    def _rpc_failed(self, w: str, e: grpc.RpcError, phase: str):
        code = e.code() if hasattr(e, "code") else None
        if code == grpc.StatusCode.FAILED_PRECONDITION:
            self.deposed = True            # a newer coordinator owns the workers: stand down
            self.stop_event.set()
            log.warning("deposed by worker %s during %s: %s", w, phase, e.details())
            return
        if code == grpc.StatusCode.ABORTED:
            log.warning("worker %s: %s aborted (%s); skipped this round", w, phase, e.details())
            if "aggregation failed" in (e.details() or ""):
                self._need_reset = True
            return
        self.members.mark_dead(w, f"{phase}: {code}")
        self._need_reset = True

    def _train_one(self, rank: int, w: str, world: int, live_ranks=None):
        md = self._metadata()
        if live_ranks is not None:
            md = md + desc.live_metadata(live_ranks)
        try:
            call = self.members.stub(w).StartTrain.with_call(
                proto.TrainRequest(rank=rank, world=world), timeout=self.cfg.train_timeout_s,
                metadata=md)
        except grpc.RpcError as e:
            self._rpc_failed(w, e, "StartTrain")
            return None
        reply, info = call
        samples = None
        for k, v in (info.trailing_metadata() or ()):
            if k == "fedb200-samples":
                samples = int(v)
        if desc.is_descriptor(reply.message):
            return {"worker": w, "rank": rank, "descriptor": desc.decode(reply.message), "samples": samples}
        raw = ckpt.decode_wire(reply.message)
        if self.cfg.keep_uploads:
            with open(self._path(f"test_{rank}.pth"), "wb") as f:   # reference src/server.py:56-58
                f.write(raw)
        return {"worker": w, "rank": rank, "state": ckpt.load_checkpoint(raw)["net"], "samples": samples,
                "bytes": len(reply.message)}

    def _send_model_to(self, w: str, payload=None) -> bool:
        payload = payload if payload is not None else self._global_payload()
        try:
            self.members.stub(w).SendModel(proto.SendModelRequest(model=payload),
                                           timeout=self.cfg.send_timeout_s, metadata=self._metadata())
            return True
        except grpc.RpcError as e:
            self._rpc_failed(w, e, "SendModel")
            return False

    def _global_payload(self):
        return ckpt.encode_wire(ckpt.dumps(self.global_state, acc=1, epoch=1))

    def _replicate(self, payload, round_id: int):
        """K-REPL on the legacy plane: same SendModel the reference uses (src/server.py:141-142),
        plus (term, round) metadata so the backup knows which round it holds."""
        if self._backup_channel is None:
            return
        try:
            proto.TrainerStub(self._backup_channel).SendModel(
                proto.SendModelRequest(model=payload), timeout=self.cfg.send_timeout_s,
                metadata=desc.fence_metadata(self.term, round_id))
            self.backup_up = True
        except grpc.RpcError as e:
            self.backup_up = False          # B5: only the backup flag changes
            log.warning("backup push failed: %s", e.code() if hasattr(e, "code") else e)

    def aggregate(self, uploads: List[dict]) -> OrderedDict:
        """Legacy-plane FedAvg (C5).  The result is what workers load: int64 counters truncated
        (N3), everything else fp32."""
        sds = [u["state"] for u in uploads]
        weights = None
        if self.cfg.weighted and all(u.get("samples") for u in uploads):
            weights = [u["samples"] for u in uploads]
        avg = fedavg_state_dicts(sds, weights)
        out = OrderedDict()
        for k, v in avg.items():
            out[k] = v.trunc().to(torch.int64) if not sds[0][k].dtype.is_floating_point else v
        return out

    # ------------------------------------------------------------------ the round loop (C4)
    def run_round(self) -> Optional[dict]:
        t0 = time.perf_counter()
        live = self.members.live_workers()
        world = len(self.cfg.workers)
        faults.fire("round_start", self.round)
        if not live:
            log.warning("round %d: no live workers", self.round)
            time.sleep(self.cfg.membership_period_s)
            return None
        peer = self.cfg.plane == "peer"
        live_ranks = [list(self.cfg.workers).index(w) for w in live] if peer else None
        if peer and self._need_reset:
            # a previous round lost a worker inside the flag protocol: put every survivor on a fresh epoch
            reset = desc.encode("reset", epoch=(self.round + 1) * 1024 + self.term)
            with ThreadPoolExecutor(max_workers=len(live)) as pool:
                list(pool.map(lambda w: self._send_model_to(w, reset), live))
            self._need_reset = False
            live = self.members.live_workers()
            live_ranks = [list(self.cfg.workers).index(w) for w in live]
            if not live:
                return None
        with ThreadPoolExecutor(max_workers=len(live)) as pool:
            uploads = list(pool.map(lambda rw: self._train_one(rw[0], rw[1], world, live_ranks), enumerate(live)))
        uploads = [u for u in uploads if u is not None]
        if self.stop_event.is_set():
            return None
        t1 = time.perf_counter()
        faults.fire("after_train", self.round)
        if not uploads:
            return None
        if peer:
            if len(uploads) != len(live):
                # somebody never trained, so the K-AGG launches already in flight on the survivors will time
                # out in their flag wait: let SendModel collect that, then reset
                self._need_reset = True
            payload = desc.encode("global", round=self.round, term=self.term, version=self.round + 1)
        else:
            self.global_state = self.aggregate(uploads)
            payload = self._global_payload()
            self.persist_model()
        t2 = time.perf_counter()
        faults.fire("after_agg", self.round)
        if self._repl_thread is not None:
            self._repl_thread.join()
        if self.cfg.replicate == "sync":
            self._replicate(payload, self.round)
        else:
            self._repl_thread = threading.Thread(target=self._replicate, args=(payload, self.round), daemon=True)
            self._repl_thread.start()
        t3 = time.perf_counter()
        faults.fire("after_repl", self.round)
        targets = [u["worker"] for u in uploads if self.members.snapshot().get(u["worker"])]
        with ThreadPoolExecutor(max_workers=max(len(targets), 1)) as pool:
            sent = list(pool.map(lambda w: self._send_model_to(w, payload), targets))
        if peer and not all(sent):
            # the aggregation did not complete on every participant (a peer died inside the flag protocol):
            # the round is NOT committed; survivors are reset and the same round number runs again
            self._need_reset = True
            return None
        t4 = time.perf_counter()
        faults.fire("round_end", self.round)
        if peer and self.cfg.checkpoint_every and (self.round + 1) % self.cfg.checkpoint_every == 0 and targets:
            self._send_model_to(targets[0], desc.encode("dump", path=os.path.abspath(self._path(GLOBAL_MODEL_FILE))))
        rec = self.sink.write(
            "round", round=self.round, term=self.term, live=len(live), uploads=len(uploads), plane=self.cfg.plane,
            world=world, train_s=t1 - t0, agg_s=t2 - t1, repl_s=t3 - t2, bcast_s=t4 - t3,
            total_s=t4 - t0, bytes_up=sum(u.get("bytes", 0) for u in uploads),
            bytes_down=len(payload) * (len(targets) + (1 if self._backup_channel else 0)))
        self.round_times.append(t4 - t0)
        return rec

    def run(self, rounds: Optional[int] = None) -> int:
        """Runs rounds ``self.round .. rounds-1``; returns the number of committed rounds."""
        rounds = self.cfg.rounds if rounds is None else rounds
        self.start_membership()
        done = 0
        if self.global_state is not None and self.round > 0 and self.cfg.plane != "peer":
            # promotion / restart: put every worker back on the last committed model (B6)
            for w in self.members.live_workers():
                self._send_model_to(w)
        while self.round < rounds and not self.stop_event.is_set():
            log.info("term %d: starting round %d", self.term, self.round)
            rec = self.run_round()
            if rec is not None:
                done += 1
                self.round += 1
                self.persist_meta()
        if self._repl_thread is not None:
            self._repl_thread.join()
        return done

    def stop(self):
        self.stop_event.set()

    def close(self):
        self.stop()
        if self._member_thread is not None:
            self._member_thread.join(timeout=self.cfg.membership_period_s + 3)
        self.members.close()
        if self._backup_channel is not None:
            self._backup_channel.close()
