"""Training worker: the gRPC *server* that a coordinator dials.

RPC behaviour kept from the reference (src/client.py:15-52):
* ``StartTrain(rank, world)``  one local pass over shard (rank, world), reply = the trained model;
* ``SendModel(model)``         install the global model, then evaluate on the whole test set
                               inside the RPC (N8), reply "success";
* ``HeartBeat``                status = 1;
* ``CheckIfPrimaryUp``         not implemented (UNIMPLEMENTED), like the reference worker.

Fixed on purpose:
* B2  a worker runs ONE pass at a time: a second ``StartTrain`` while one is running is refused
      with ABORTED instead of racing on the shared model;
* fencing  requests that carry a term older than the newest term seen are refused with
      FAILED_PRECONDITION, so a deposed coordinator cannot drive the worker (no split brain).
      Reference coordinators send no term and are always accepted;
* B7/B8  nothing happens at import, the checkpoint directory is created on demand.

Planes: ``legacy`` answers with base64 checkpoints (works with a reference coordinator);
``peer`` answers with a descriptor - the weights are already visible to the coordinator through
peer-mapped GPU memory and the broadcast lands in this worker's arena without an RPC payload.
"""
from __future__ import annotations

import os
import threading
from dataclasses import dataclass
from typing import Optional

import grpc

from .. import proto
from ..utils import checkpoint as ckpt
from ..utils.metrics import JsonlSink, log
from . import descriptor as desc
from .trainer import LocalTrainer


@dataclass
class WorkerConfig:
    address: str = "temp"                 # listen address AND checkpoint stem (reference -a)
    compress: bool = False                # reference -c Y
    plane: str = "legacy"                 # legacy | peer
    eval_on_send: bool = True             # N8: evaluate inside SendModel
    write_checkpoint: bool = True         # keep ./checkpoint/<address>.pth current
    checkpoint_root: str = "./checkpoint"
    dump_root: str = "."                  # "dump" descriptors may only write below this directory
    max_rpc_workers: int = 10


class WorkerServicer(proto.TrainerServicer):
    def __init__(self, trainer: LocalTrainer, cfg: WorkerConfig, sink: Optional[JsonlSink] = None,
                 peer_endpoint=None):
        self.trainer, self.cfg = trainer, cfg
        self.sink = sink or JsonlSink(None)
        self.peer = peer_endpoint            # fedb200.parallel.mem.PeerEndpoint on the GPU plane
        self._busy = threading.Lock()
        self._term_lock = threading.Lock()
        self.term = 0
        self.round = -1
        self.passes = 0
        self.rejected = 0
        self.last_eval = None
        if cfg.write_checkpoint:             # initial checkpoint, as the reference writes at import
            self._persist()

    # ------------------------------------------------------------------ helpers
    def _ckpt_path(self):
        return ckpt.worker_checkpoint_path(self.cfg.address, self.cfg.checkpoint_root)

    def _persist(self):
        self.trainer.save(self._ckpt_path(), epoch=1)

    def _fence(self, context) -> bool:
        term, rnd = desc.read_fence(context.invocation_metadata())
        if term is None:
            return True                      # reference coordinator: no fencing data
        with self._term_lock:
            if term < self.term:
                self.rejected += 1
                context.abort(grpc.StatusCode.FAILED_PRECONDITION,
                              f"stale term {term} < {self.term}")
            self.term = term
            if rnd is not None:
                self.round = rnd
        return True

    # ------------------------------------------------------------------ rpcs
    def StartTrain(self, request, context):
        self._fence(context)
        if not self._busy.acquire(blocking=False):
            self.rejected += 1
            context.abort(grpc.StatusCode.ABORTED, "a local pass is already running")
        try:
            if self.cfg.plane == "peer" and self.peer is not None:
                st = self.peer.start_round(request.rank, request.world, desc.read_live(context.invocation_metadata()),
                                           round_id=self.round)
            else:
                st = self.trainer.train(1, request.rank, request.world)
            self.passes += 1
            self.sink.write("train_pass", address=self.cfg.address, rank=request.rank,
                            world=request.world, steps=st.steps, samples=st.samples,
                            loss=st.mean_loss, acc=st.accuracy, seconds=st.seconds,
                            term=self.term, round=self.round)
            if self.cfg.plane == "peer":
                msg = desc.encode("upload", samples=st.samples, steps=st.steps, loss=st.mean_loss,
                                  correct=st.correct, arena=self.trainer.arena.manifest.fingerprint(),
                                  endpoint=(self.peer.name if self.peer else None),
                                  round=self.round, term=self.term)
                return proto.TrainReply(message=msg)
            if self.cfg.write_checkpoint:
                self._persist()
            context.set_trailing_metadata((("fedb200-samples", str(st.samples)),))
            return proto.TrainReply(message=ckpt.encode_wire(self.trainer.dumps(epoch=1)))
        finally:
            self._busy.release()

    def SendModel(self, request, context):
        self._fence(context)
        with self._busy:                     # never swap weights under a running pass
            if desc.is_descriptor(request.model):
                d = desc.decode(request.model)
                kind = d.get("kind")
                if self.peer is None:
                    context.abort(grpc.StatusCode.FAILED_PRECONDITION, "descriptor received but this worker has no GPU plane")
                if kind == "reset":
                    self.peer.reset(int(d["epoch"]))
                    return proto.SendModelReply(reply="success")
                if kind == "remap":
                    self.peer.remap(int(d["rank"]))
                    return proto.SendModelReply(reply="success")
                if kind == "pull":
                    self.peer.pull_global(int(d["src"]) if "src" in d else None)
                    return proto.SendModelReply(reply="success")
                if kind == "dump":
                    # the path comes in over an unauthenticated channel: only files below the configured root
                    root = os.path.realpath(self.cfg.dump_root)
                    path = os.path.realpath(d["path"])
                    if os.path.commonpath([root, path]) != root:
                        context.abort(grpc.StatusCode.PERMISSION_DENIED, f"dump path outside {root}")
                    self.peer.dump(path, what=d.get("what", "global"))
                    return proto.SendModelReply(reply="success")
                if kind == "status":
                    self.sink.write("peer_status", address=self.cfg.address, round=self.round, term=self.term,
                                    **self.peer.status())
                    return proto.SendModelReply(reply="success")
                if kind != "global":
                    context.abort(grpc.StatusCode.INVALID_ARGUMENT, f"unexpected descriptor {d}")
                try:
                    ev = self.peer.finish_round(evaluate=self.cfg.eval_on_send)
                except Exception as e:   # noqa: BLE001  (FedbError: the flag protocol gave up on a dead peer)
                    self.sink.write("aggregation_failed", address=self.cfg.address, error=str(e), round=self.round)
                    context.abort(grpc.StatusCode.ABORTED, f"aggregation failed: {e}")
                prof = self.peer.take_profile() if hasattr(self.peer, "take_profile") else {}
                if prof:                 # --profile: device-timed phases of this round on this worker's GPU
                    self.sink.write("round_profile", address=self.cfg.address, round=self.round, term=self.term, **prof)
                if ev is not None:
                    self.last_eval = ev
                    self.sink.write("eval", address=self.cfg.address, loss=ev.mean_loss, acc=ev.accuracy,
                                    samples=ev.samples, seconds=ev.seconds, term=self.term, round=self.round)
                return proto.SendModelReply(reply="success")
            else:
                state = ckpt.load_checkpoint(ckpt.decode_wire(request.model))
                self.trainer.load_state_dict(state["net"])
            if self.cfg.write_checkpoint:
                self._persist()
            if self.cfg.eval_on_send:
                ev = self.trainer.test()
                self.last_eval = ev
                self.sink.write("eval", address=self.cfg.address, loss=ev.mean_loss, acc=ev.accuracy,
                                samples=ev.samples, seconds=ev.seconds, term=self.term, round=self.round)
        return proto.SendModelReply(reply="success")

    def HeartBeat(self, request, context):
        return proto.HeartBeatResponse(status=1)


def serve_worker(trainer: LocalTrainer, cfg: WorkerConfig, sink=None, peer_endpoint=None,
                 block: bool = True):
    """Starts the worker's gRPC server on ``cfg.address``; returns (server, servicer)."""
    server = proto.make_server(cfg.max_rpc_workers, cfg.compress)
    servicer = WorkerServicer(trainer, cfg, sink, peer_endpoint)
    proto.add_TrainerServicer_to_server(servicer, server)
    bound = server.add_insecure_port(cfg.address)
    if bound == 0:
        raise OSError(f"cannot bind worker address {cfg.address!r}")
    server.start()
    log.info("worker listening on %s (plane=%s compress=%s)", cfg.address, cfg.plane, cfg.compress)
    if block:
        server.wait_for_termination()
    return server, servicer
