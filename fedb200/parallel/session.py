"""``GpuFederation`` - the public, in-process API of the GPU data plane (what bench.py times).

One instance per rank (process).  ``run_round()`` is a whole federated round of the reference
(src/server.py:120-153 + src/client.py:16-31) for this rank:

    H2D of the shard's uint8 images/labels (pinned) -> local pass over shard (rank, world) [N9] ->
    K-AGG / K-BCAST (/K-REPL) -> evaluation of the new global model on the test set [N8] ->
    one 32-byte D2H read of the loss/accuracy counters.

The gRPC control plane (``fedb200.control``) drives exactly these calls from ``StartTrain`` /
``SendModel`` when workers run with ``--plane peer``; here they are driven directly, rank-symmetric,
which is also how a torchrun job uses the framework.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from ..control.trainer import LocalTrainer, TrainerConfig
from ..ops import _lib
from ..utils import data as D
from ..utils.metrics import DeviceTimer, device_phase
from ..ops import cuda as K
from .group import FedGroup
from .mem import PeerGroup, PeerLayout, tensor_from_ptr


def dist_min_barrier():
    """all-reduce(MIN) on a flag tensor through torch.distributed (the plumbing backend)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return None

    def fn(flag: torch.Tensor):
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return fn


class TestExchange:
    """All-gather of the evaluation images over NVLink (K-PUSH, csrc/fedavg.cu).

    Every client evaluates the global model on the SAME test set every round (N8).  Instead of N processes each
    pulling all of it through their own PCIe link, rank r uploads rows ``rows(r)`` from its pinned copy and pushes
    them into every peer's buffer - one ``multimem.st`` stream through the NVLS multicast alias, or unicast peer
    stores.  The buffer is double (one half per round parity): a rank can run at most one K-AGG ahead of its peers, so
    round r+2's pushes never land in a half that somebody still evaluates.  Ordering of the pushes before the peers'
    evaluation comes from the round's K-AGG flags: a rank raises its "arrived" flag after its own pushes (same stream
    order), and nobody leaves the aggregation before everybody has arrived.
    """
    __test__ = False

    def __init__(self, group: FedGroup, n_rows: int, row_bytes: int, tag: str, dist_barrier=None, use_multicast: bool = True):
        self.group, self.rank, self.world = group, group.rank, group.world
        self.n_rows, self.row_bytes = n_rows, row_bytes
        self.half_bytes = (n_rows * row_bytes + 255) // 256 * 256
        layout = PeerLayout(n_f32=2 * self.half_bytes // 4, n_i64=0, with_codec=False)
        self.peers = PeerGroup(self.rank, self.world, group.device, layout, tag, use_multicast=use_multicast,
                               barrier=dist_barrier)
        me = self.peers.me
        self._views = [tensor_from_ptr(me.client_f32_ptr() + h * self.half_bytes, n_rows * row_bytes, torch.uint8,
                                       group.device) for h in (0, 1)]

    def enabled(self) -> bool:
        """Only while every rank takes part: a dead rank's slice would never arrive (then each rank uploads all rows)."""
        return len(self.group.live) == self.world

    def rows(self, r: int):
        return r * self.n_rows // self.world, (r + 1) * self.n_rows // self.world

    def slice_bytes(self) -> int:
        lo, hi = self.rows(self.rank)
        return (hi - lo) * self.row_bytes

    def images(self, half: int) -> torch.Tensor:
        return self._views[half].view(self.n_rows, *self._shape)

    def set_shape(self, shape) -> None:
        self._shape = tuple(shape)

    def push(self, half: int, stream: int) -> None:
        lo, hi = self.rows(self.rank)
        off = half * self.half_bytes + lo * self.row_bytes
        nbytes = (hi - lo) * self.row_bytes
        bufs = self.peers.buffers
        mc = self.peers.mc_client_f32_ptr()
        if mc:
            # the multicast store also lands in my own copy, which already holds the bytes: harmless
            K.push_copy(bufs[self.rank].client_f32_ptr() + off, [], nbytes, mc_dst=mc + off, stream=stream)
        else:
            K.push_copy(bufs[self.rank].client_f32_ptr() + off,
                        [bufs[r].client_f32_ptr() + off for r in range(self.world) if r != self.rank], nbytes, stream=stream)

    def close(self):
        self.peers.close()


class GpuFederation:
    def __init__(self, model: str = "LeNet", rank: int = 0, world: int = 1, device: Optional[int] = None,
                 dataset: Optional[D.ImageDataset] = None, seed: Optional[int] = None, tag: Optional[str] = None,
                 lr: float = 0.1, batch_size: int = 128, augment: bool = False, evaluate: bool = True,
                 test_allgather: Optional[bool] = None, eval_mode: str = "inline", profile: bool = False, **group_kw):
        """``eval_mode``: where the evaluation of a round's global model runs (every client scores the whole test set
        every round in all modes, N8).  "inline" (default): on the round's critical path, as the reference does
        (src/client.py:30).  "overlap": on a second stream, from a snapshot of the global model, concurrently with the
        next round's local pass - ``run_round_async`` then reports it one round late.  Measured on B200 (LeNet,
        profiles/README.md): while ANY kernel of another stream is resident, every kernel boundary of the training
        stream costs about 3 us more, and a LeNet pass has two boundaries per 31 us step - the overlap loses more than
        the 0.42 ms evaluation it hides (79.1 -> 74.8 rounds/s at N = 1), so it is opt-in, for models whose kernels are
        long enough not to care."""
        device = rank if device is None else device
        torch.cuda.set_device(device)
        if eval_mode not in ("overlap", "inline"):
            raise ValueError(f"eval_mode {eval_mode!r} (overlap | inline)")
        self.rank, self.world, self.evaluate = rank, world, evaluate
        self.eval_mode = eval_mode if evaluate else "inline"
        ds = dataset if dataset is not None else D.synthetic_cifar(seed=0)
        cfg = TrainerConfig(model=model, lr=lr, batch_size=batch_size, device=f"cuda:{device}",
                            seed=(rank if seed is None else seed), augment=augment)
        self.trainer = LocalTrainer(cfg, ds.pin())
        self.engine = self.trainer._engine
        tag = tag or os.environ.get("MASTER_PORT", "solo") + f"-{os.getppid()}"
        self.group = FedGroup(self.trainer, rank, world, tag, dist_barrier=dist_min_barrier(), **group_kw)
        self.rounds = 0
        # profile=True: NVTX ranges + CUDA-event pairs around the phases of every round (train / agg / repl / repl_wait /
        # eval); take_profile() turns them into one record (device milliseconds, NVLink bytes, GB/s).  Costs a device
        # synchronisation per record, so it is off in timed runs.
        self.timer = DeviceTimer() if profile else None
        self.group.timer = self.timer
        if self.eval_mode == "overlap":
            self.engine.enable_snapshot_eval()
        # evaluation-set all-gather over NVLink: on by default for rank-symmetric multi-GPU sessions (torchrun); workers
        # that may be restarted one by one (gRPC plane: persistent_server / rejoin) keep the plain per-rank upload
        if test_allgather is None:
            test_allgather = world > 1 and evaluate and not (group_kw.get("persistent_server") or group_kw.get("rejoin"))
        self.exchange = None
        if test_allgather and world > 1:
            tx = ds.test_x
            row = int(tx[0].numel()) * tx.element_size()
            if row % 16 == 0:
                self.exchange = TestExchange(self.group, int(tx.shape[0]), row, tag + "-tst", dist_barrier=dist_min_barrier(),
                                             use_multicast=group_kw.get("use_multicast", True))
                self.exchange.set_shape(tx.shape[1:])
                self.engine.set_test_exchange(self.exchange)

    def run_round(self, upload: bool = True, sync: bool = True):
        """One federated round.  Returns (train_stats, eval_stats) of THIS round when ``sync`` (an overlapped evaluation is
        waited for); with ``sync=False`` everything is only enqueued (device-timed benchmarking)."""
        live = self.group.live
        batches = self.trainer.shard(live.index(self.rank), self.world) if self.rank in live else []
        self.group.before_local_pass()
        with device_phase(self.timer, "train"):
            self.engine.train_pass(batches, upload=upload, sync=False, with_test=self.evaluate)
        if self.eval_mode == "overlap":
            self._join_eval(keep=True)
        self.group.aggregate()
        if self.evaluate:
            if self.eval_mode == "overlap":
                self._launch_overlapped_eval(upload)
            else:
                with device_phase(self.timer, "eval"):
                    self.engine.eval_pass(upload=upload, sync=False)
        self.rounds += 1
        if not sync:
            return None
        if self.eval_mode == "overlap":
            self._join_eval()                # a synchronous caller wants THIS round's score: finish it now
        out = self.engine.read_round_stats()
        self.group.check()
        return out

    # ---- evaluation on a second stream (eval_mode="overlap")
    def _join_eval(self, keep: bool = False) -> None:
        """See ``CudaEngine.join_overlapped_eval``: a rank only announces itself to K-AGG after this point."""
        self.engine.join_overlapped_eval(keep=keep)

    def _launch_overlapped_eval(self, upload: bool) -> None:
        eng = self.engine
        cur = torch.cuda.current_stream(eng.dev)
        eng.snapshot()                                   # the new global model, before the next pass trains on it
        if getattr(eng, "_eval_done", None) is None:
            eng._eval_done, self._snap_ready = torch.cuda.Event(), torch.cuda.Event()
        self._snap_ready.record(cur)
        with torch.cuda.stream(eng.eval_stream):
            eng.eval_stream.wait_event(self._snap_ready)
            with device_phase(self.timer, "eval", eng.eval_stream):
                if eng.queue_eval_available(upload):
                    eng.queue_eval_start()               # LeNet: background CTAs on the SMs the training step leaves idle
                else:
                    eng.eval_pass(upload=upload, sync=False, snapshot=True)
            eng._eval_done.record(eng.eval_stream)

    def take_profile(self) -> dict:
        """Device-timed phases since the last call (synchronises): milliseconds per phase, the NVLink bytes of this rank's
        last aggregation and the rate they were moved at, and how much of the replication was hidden."""
        if self.timer is None:
            return {}
        ms = self.timer.ms(reset=True)
        out = {f"{k}_ms": round(v, 4) for k, v in ms.items()}
        g = self.group
        out.update(rank=self.rank, rounds=self.rounds, live=len(g.live), nvlink_bytes=g.nvlink_bytes_last,
                   arena_bytes=g.man.n_f32 * 4)
        if ms.get("agg"):
            out["agg_gbps"] = round(g.nvlink_bytes_last / (ms["agg"] * 1e-3) / 1e9, 2)
        if ms.get("repl"):
            out["repl_hidden_fraction"] = round(max(0.0, 1.0 - ms.get("repl_wait", 0.0) / ms["repl"]), 4)
        return out

    def drain(self, sync: bool = True):
        """Waits (on the stream; on the host too when ``sync``) for the evaluation that is still in flight and returns
        its statistics: the score of the LAST round's global model."""
        if self.eval_mode == "overlap":
            self._join_eval(keep=True)
        if not sync:
            return None
        torch.cuda.current_stream(self.engine.dev).synchronize()
        return self.engine.read_stats("eval")

    def run_round_async(self, upload: bool = True) -> "RoundHandle":
        """One federated round, enqueued only.  The round's loss / accuracy counters and the K-AGG status travel to a
        pinned slot at the end of the round; ``handle.result()`` waits for exactly that copy.  Calling it after the
        NEXT round has been enqueued keeps the GPU busy across the round boundary (up to three rounds may be in flight).
        With ``eval_mode="overlap"`` the evaluation half of the result is the one that finished last, i.e. the score of
        the PREVIOUS round's global model (``handle.eval_round``; zero samples for the first round); ``drain()``
        returns the last one."""
        self.run_round(upload=upload, sync=False)
        c = self.group.peers.me.ctrl()
        off = _lib.CTRL_OFF["status"] // 4
        lag = self.eval_mode == "overlap"
        h = RoundHandle(self, self.engine.enqueue_round_stats(c[off:off + 1],
                                                              eval_src=self.engine.eval_stats_prev if lag else None))
        h.round, h.eval_round = self.rounds, self.rounds - 1 if lag else self.rounds
        return h

    def aggregation_selfcheck(self, all_gather=None) -> dict:
        """Proof that the multi-GPU aggregation did what the reference's allreduce() does (src/server.py:155-179), for
        hosts that cannot look into the GPUs: one extra local pass, the pre-aggregation arenas of ALL ranks are gathered
        (``all_gather(tensor) -> list of tensors``, e.g. over torch.distributed), K-AGG runs, and (i) the result is
        compared bit for bit with the CPU oracle fed with the gathered arenas, (ii) a 64-bit checksum of every rank's
        fp32 + int64 arena is compared across ranks.  Lossless codec only (fp8 deltas are not bit-exact by design)."""
        from .fedavg import fedavg_flat_cpu
        man = self.group.man
        live = self.group.live
        batches = self.trainer.shard(live.index(self.rank), self.world)
        self.group.before_local_pass()
        self.engine.train_pass(batches, upload=False, sync=False, with_test=False)
        if self.eval_mode == "overlap":
            self._join_eval()
        torch.cuda.synchronize()
        a = self.trainer.arena
        pre32, pre64 = a.f32[:man.n_f32].clone(), a.i64[:man.n_i64].clone()
        gather = all_gather if all_gather is not None else (lambda t: [t])
        all32 = [t.cpu() for t in gather(pre32)]
        all64 = [t.cpu() for t in gather(pre64)] if man.n_i64 else [pre64.cpu() for _ in all32]
        self.group.aggregate()
        torch.cuda.synchronize()
        self.group.check()
        post32, post64 = a.f32[:man.n_f32].clone(), a.i64[:man.n_i64].clone()
        want32, want64 = torch.empty(man.n_f32), torch.empty(man.n_i64, dtype=torch.int64)
        fedavg_flat_cpu(all32, all64, None, want32, want64)
        exact = torch.equal(post32.cpu(), want32) and torch.equal(post64.cpu(), want64)
        chk = (post32.view(torch.int32).to(torch.int64).sum() + post64.sum()).reshape(1)
        sums = [int(t) for t in gather(chk)]
        return {"agg_check": "bit-exact" if exact else "failed", "arenas_equal": len(set(sums)) == 1,
                "ranks_gathered": len(all32), "checksum": sums[0]}

    def h2d_bytes_per_round(self) -> int:
        return self.engine.h2d_bytes_last

    def d2h_bytes_per_round(self) -> int:
        return 32 + 4

    def close(self):
        if self.exchange is not None:
            torch.cuda.synchronize()
            self.exchange.close()
        self.group.close()


class RoundHandle:
    def __init__(self, fed: GpuFederation, ticket):
        self.fed, self.ticket = fed, ticket

    def result(self):
        train, ev, status = self.fed.engine.fetch_round_stats(self.ticket)
        if status:
            raise _lib.FedbError(f"K-AGG flag wait failed on rank {self.fed.rank}: status {status} "
                                 "(1 = a peer never arrived, 2 = a peer never finished, 3 = aborted)")
        return train, ev


class PeerWorker:
    """What ``WorkerServicer`` drives on the GPU plane: ``StartTrain`` -> ``start_round`` (local pass, then the
    K-AGG launch is enqueued and overlaps the RPC round trip), ``SendModel`` -> ``finish_round`` (wait for the
    aggregation, verify the flag protocol, evaluate).  gRPC handler threads differ per call, so every entry
    point re-selects the device."""

    def __init__(self, fed: GpuFederation):
        self.fed = fed
        self.name = f"rank{fed.rank}"
        self.device = fed.trainer.device

    def start_round(self, rank: int, world: int, live=None, round_id=None):
        torch.cuda.set_device(self.device)
        fed = self.fed
        if live is not None:
            fed.group.set_live(live)
        if round_id is not None:
            # the replica's version stamp is the COORDINATOR's round number (+1 = rounds committed once this one is),
            # not a per-rank launch count: a round that failed and was re-run must not advance it twice
            fed.group.round = int(round_id)
        batches = fed.trainer.shard(rank, world)
        fed.group.before_local_pass()
        with device_phase(fed.timer, "train"):
            st = fed.engine.train_pass(batches, upload=True, sync=True, with_test=True)
        fed.group.aggregate()
        return st

    def finish_round(self, evaluate: bool = True):
        torch.cuda.set_device(self.device)
        torch.cuda.current_stream(self.device).synchronize()
        self.fed.group.side_stream.synchronize()     # holder: global model + replica of this round are complete
        self.fed.group.check()                       # raises FedbError when a peer never showed up
        if not evaluate:
            return None
        with device_phase(self.fed.timer, "eval"):
            return self.fed.engine.eval_pass(upload=True, sync=True)

    def take_profile(self) -> dict:
        return self.fed.take_profile()

    def reset(self, epoch: int) -> None:
        torch.cuda.set_device(self.device)
        self.fed.group.reset_sync(epoch)

    def remap(self, r: int) -> None:
        torch.cuda.set_device(self.device)
        self.fed.group.remap_peer(r)

    def pull_global(self, src=None) -> None:
        torch.cuda.set_device(self.device)
        self.fed.group.pull_global(src)

    def dump(self, path: str, what: str = "global") -> None:
        """Writes the server's global model (aux arena of the current holder) or, ``what="client"``, this worker's own
        arena as a reference-format checkpoint."""
        from ..utils import checkpoint as ckpt
        torch.cuda.set_device(self.device)
        torch.cuda.synchronize(self.device)
        arena = self.fed.group.global_arena if what == "global" else self.fed.trainer.arena
        ckpt.save_checkpoint(path, arena.state_dict(to_cpu=True), acc=1, epoch=1)

    dump_global = dump

    def status(self) -> dict:
        """Roles, replica version and a checksum of this worker's arena (one small D2H read): what an operator - or
        tools/grpc_gpu_check.py - needs to see that the data plane followed a membership change."""
        torch.cuda.set_device(self.device)
        torch.cuda.synchronize(self.device)
        g = self.fed.group
        a = self.fed.trainer.arena
        chk = int(a.f32[:g.man.n_f32].view(torch.int32).to(torch.int64).sum()) + (int(a.i64.sum()) if g.man.n_i64 else 0)
        return {"rank": g.rank, "primary_rank": g.primary_rank, "backup_rank": g.backup_rank, "live": list(g.live),
                "replica_version": K.ctrl_field(g.peers.me.ctrl(), "version"), "arena_checksum": chk,
                "multicast": bool(g.peers.mc_base) and not g._remapped, "agg_round": g.round}
