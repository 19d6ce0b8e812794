"""GPU data plane of a federation on one NVSwitch box: one process per GPU, every worker's arena
peer-mapped, FedAvg + broadcast (+ replica) as ONE kernel launch per rank per round.

What the reference does per round with 2N+1 RPCs carrying base64 checkpoints and a CPU mean over
files (src/server.py:113-179) happens here as:

    train pass (CUDA graph) -> K-AGG two-shot (csrc/fedavg.cu) -> [K-REPL on a side stream] -> eval pass

* every rank owns the shard ``split_ranges(n_f32, live)[i]`` of the flat arena: it pulls that shard from
  every live peer over NVLink, reduces it in the reference's order, and pushes the result into every
  peer's arena, the primary's global model and (fused mode) the backup's replica;
* the device-flag protocol inside the kernel replaces both thread-join barriers; spins are bounded and
  surface as ``FedbError`` instead of a hang when a peer died;
* with NVLS the push is one ``multimem.st`` and (opt-in, order not bit-reproducible) the pull is one
  ``multimem.ld_reduce``;
* ``codec="fp8"`` ships block-scaled e4m3 deltas with error feedback instead of fp32 words (K-CODEC).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from ..ops import _lib
from ..ops import cuda as K
from ..utils.metrics import device_phase
from .arena import Arena, assign_roles, split_ranges
from .mem import PeerGroup, PeerLayout


class FedGroup:
    def __init__(self, trainer, rank: int, world: int, tag: str, *, use_multicast: bool = True,
                 nvls_reduce: bool = False, codec: str = "none", primary_rank: int = 0,
                 backup_rank: Optional[int] = None, repl_mode: str = "overlap", dist_barrier=None,
                 timeout_ns: int = K.DEFAULT_TIMEOUT_NS, persistent_server: bool = False, rejoin: bool = False):
        assert trainer.device.type == "cuda"
        self.trainer, self.rank, self.world = trainer, rank, world
        self.device = trainer.device.index or 0
        self.man = trainer.arena.manifest
        self.codec, self.nvls_reduce, self.repl_mode = codec, nvls_reduce, repl_mode
        # roles follow the membership: the preferred ranks hold the global model / the replica while they are alive,
        # otherwise the two lowest live ranks do (the reference re-admits ANY worker, src/server.py:78-101, so no rank
        # may be special); see _assign_roles
        self._pref_primary = primary_rank
        self._pref_backup = (primary_rank + 1) % world if backup_rank is None else backup_rank
        self.primary_rank, self.backup_rank = self._pref_primary, self._pref_backup
        self.timeout_ns = timeout_ns
        self.layout = PeerLayout.for_manifest(self.man, with_codec=True)
        torch.cuda.set_device(self.device)
        self.peers = PeerGroup(rank, world, self.device, self.layout, tag, use_multicast=use_multicast,
                               barrier=dist_barrier, persistent_server=persistent_server, rejoin=rejoin)
        # move the trainer's live state into the peer-visible allocation
        old = trainer.arena
        new = self.peers.me.client_arena(self.man)
        new.copy_from(old)
        new.grad, new.momentum, new.sgd_steps = old.grad, old.momentum, old.sgd_steps
        new.bind(trainer.model)
        trainer.arena = new
        if trainer._engine is not None:
            trainer._engine.rebind_arena(new)
        self.live: List[int] = list(range(world))
        self.round = 0
        self.side_stream = torch.cuda.Stream(device=self.device, priority=0)
        self._agg_done = torch.cuda.Event()
        self._repl_done = torch.cuda.Event()
        self._side_ctrl = K.new_ctrl(trainer.device)      # completion counter of the side-stream copy kernel
        self._common_base = False
        self._repl_pending = False
        self.timer = None                 # DeviceTimer when the owner profiles (GpuFederation(profile=True))
        self.nvlink_bytes_last = 0        # algorithmic bytes this rank moved over NVLink in its last K-AGG (+ K-REPL)
        self._wait_ev = None
        self._remapped = False            # a restarted peer's new memory is not bound to the multicast object
        if codec == "fp8":
            dev = trainer.device
            self.residual = torch.zeros(self.man.n_f32, dtype=torch.float32, device=dev)
            self.base = torch.zeros(self.man.n_f32, dtype=torch.float32, device=dev)   # last global model
            self.base.copy_(new.f32[:self.man.n_f32])
        torch.cuda.synchronize(self.device)

    # ------------------------------------------------------------------ views
    @property
    def global_arena(self) -> Arena:
        """The server's global model (lives on the primary's GPU, visible everywhere)."""
        return self.peers.buffers[self.primary_rank].aux_arena(self.man)

    @property
    def replica_arena(self) -> Arena:
        return self.peers.buffers[self.backup_rank].aux_arena(self.man)

    def set_live(self, live: Sequence[int]) -> None:
        """Membership change: only these ranks take part in (and are averaged by) the next rounds."""
        self.live = sorted(live)
        self._assign_roles()

    def _assign_roles(self) -> None:
        """Who keeps the server's global model (aux arena of ``primary_rank``) and its replica (``backup_rank``): the
        preferred ranks while they are live, else the lowest live ranks.  Every rank derives the same answer from the
        same live list, and the holder refreshes both copies from its own client arena right after each K-AGG, so a
        newly appointed holder is up to date after the first committed round of its tenure."""
        if self.live:
            self.primary_rank, self.backup_rank = assign_roles(self.live, self._pref_primary, self._pref_backup)

    # ------------------------------------------------------------------ one aggregation
    def aggregate(self, weights: Optional[Sequence[float]] = None) -> None:
        """K-AGG launch of THIS rank (every live rank calls it once per round, any time after its own
        training pass was enqueued)."""
        if self.rank not in self.live:
            return
        bufs = self.peers.buffers
        live = self.live
        me = live.index(self.rank)
        begin, end = split_ranges(self.man.n_f32, len(live))[me]
        # NVLink budget: every rank already RECEIVES (N-1) shards as reads and (N-1) shards as broadcast writes.
        # The server's global model is therefore not written by the peers (that would add (N-1) more inbound
        # shards on the primary's GPU): after K-AGG the primary's own client arena IS the global model, and it
        # copies it into the aux arena locally (below).  Only the opt-in "fused" mode writes the replica here.
        fused_repl = self.repl_mode == "fused" and self.backup_rank != self.primary_rank
        dst, dst_i64 = [], []
        if fused_repl:
            dst.append(bufs[self.backup_rank].aux_f32_ptr())
            dst_i64.append(bufs[self.backup_rank].aux_i64_ptr())
        mc_dst = self.peers.mc_client_f32_ptr() if (len(live) == self.world and not self._remapped) else 0
        if not mc_dst:
            dst += [bufs[r].client_f32_ptr() for r in live]
        dst_i64 += [bufs[r].client_i64_ptr() for r in live]
        kw = {}
        # deltas need a base that is IDENTICAL on every rank: the first aggregation (different random
        # inits, N6) therefore goes out lossless and establishes it
        use_codec = self.codec == "fp8" and self._common_base
        if use_codec:
            n = self.man.n_f32
            K.codec_encode(self.trainer.arena.f32[:n], self.base, self.residual, self.peers.me.q(),
                           self.peers.me.scale())
            kw = dict(src_q=[bufs[r].q_ptr() for r in live], src_scale=[bufs[r].scale_ptr() for r in live],
                      base=self.base.data_ptr())
        elif self.nvls_reduce and mc_dst:
            kw = dict(mc_src=mc_dst)
        # bytes this rank pulls (every peer's copy of its shard) and pushes (its reduced shard to every peer; once through
        # the switch with multicast): the denominator of the achieved-GB/s figures
        shard_bytes = (end - begin) * (1 if use_codec else 4)
        n_peer = len(live) - 1
        self.nvlink_bytes_last = shard_bytes * n_peer + (end - begin) * 4 * (1 if mc_dst else n_peer + (1 if fused_repl else 0))
        with device_phase(self.timer, "agg"):
            self._launch_agg(bufs, live, me, dst, dst_i64, begin, end, weights, mc_dst, kw)
        if self.codec == "fp8":
            # the new global model is the base of the next round's deltas
            self.base.copy_(self.trainer.arena.f32[:self.man.n_f32])
            self._common_base = True
        self.round += 1
        if self.rank == self.primary_rank:
            # Server-side copy + K-REPL on a side stream, overlapped with what follows (eval, next local pass):
            # ONE kernel reads the primary's fresh client arena and writes the global model (local HBM) and, in
            # "overlap" mode, the backup's replica (peer stores over NVLink), then stamps the replica's version.
            self._agg_done.record(torch.cuda.current_stream(self.device))
            self.side_stream.wait_event(self._agg_done)
            me_buf, bk = bufs[self.primary_rank], bufs[self.backup_rank]
            to_replica = self.repl_mode == "overlap" and self.backup_rank != self.primary_rank and bk.base != 0
            d32 = [me_buf.aux_f32_ptr()] + ([bk.aux_f32_ptr()] if to_replica else [])
            d64 = [me_buf.aux_i64_ptr()] + ([bk.aux_i64_ptr()] if to_replica else [])
            if to_replica:
                self.nvlink_bytes_last += self.man.n_f32 * 4
            with torch.cuda.stream(self.side_stream), device_phase(self.timer, "repl", self.side_stream):
                K.fedavg_raw([me_buf.client_f32_ptr()], d32, 0, self.man.n_f32, ctrl_local=self._side_ctrl.data_ptr(),
                             src_i64=[me_buf.client_i64_ptr()], dst_i64=d64, n_i64=self.man.n_i64,
                             version=self.round, dst_ctrl=[0] + ([bk.ctrl_ptr] if to_replica else []),
                             gate=self.peers.me.ctrl_ptr + _lib.CTRL_OFF["status"])   # a failed K-AGG copies nothing
            self._repl_done.record(self.side_stream)
            self._repl_pending = True

    def _launch_agg(self, bufs, live, me, dst, dst_i64, begin, end, weights, mc_dst, kw) -> None:
        K.fedavg_raw([bufs[r].client_f32_ptr() for r in live], dst, begin, end,
                     ctrl_local=self.peers.me.ctrl_ptr,
                     src_i64=[bufs[r].client_i64_ptr() for r in live], dst_i64=dst_i64,
                     n_i64=self.man.n_i64 if me == 0 else 0, weights=weights, mc_dst=mc_dst,
                     sync_mode=1, rank=me, world=len(live), ctrl_peer=[bufs[r].ctrl_ptr for r in live],
                     timeout_ns=self.timeout_ns, **kw)

    def before_local_pass(self) -> None:
        """The side-stream copy READS the primary's client arena; the next local pass WRITES it.  Evaluation may
        overlap the copy, training may not: callers invoke this right before they enqueue a training pass."""
        if self._repl_pending:
            # "repl_wait" is what the replication still costs the round: the part of the side-stream copy that was NOT
            # hidden behind the evaluation / the host's work since the aggregation (hidden fraction = 1 - wait / repl)
            with device_phase(self.timer, "repl_wait"):
                torch.cuda.current_stream(self.device).wait_event(self._repl_done)
            self._repl_pending = False

    def check(self) -> None:
        """Raises if a bounded spin gave up (peer died / never arrived).  Costs one 4-byte D2H read."""
        st = K.ctrl_field(self.peers.me.ctrl(), "status")
        if st:
            raise _lib.FedbError(f"K-AGG flag wait failed on rank {self.rank}: status {st} "
                                 "(1 = a peer never arrived, 2 = a peer never finished, 3 = aborted)")

    def reset_sync(self, epoch: int) -> None:
        """After a failed round (a peer died inside the flag protocol): clear this rank's flags and jump to a
        coordinator-chosen epoch that is larger than any epoch used so far, so that stale flag values written
        by slow peers can never satisfy a future wait."""
        torch.cuda.synchronize(self.device)
        c = self.peers.me.ctrl()
        c.zero_()
        c[_lib.CTRL_OFF["epoch"] // 4] = int(epoch)
        torch.cuda.synchronize(self.device)

    def remap_peer(self, r: int) -> None:
        """A restarted worker re-joined with a fresh allocation: map it (worker re-admission)."""
        torch.cuda.synchronize(self.device)
        self.peers.remap(r)
        self._remapped = True             # from now on the fan-out uses unicast peer stores (the NVLS alias still
                                          # reaches the dead process's pages for that rank)

    def pull_global(self, src: Optional[int] = None) -> None:
        """Newcomer side of re-admission: copy the server's global model out of the holder's peer memory into
        this worker's arena (what the reference does by re-sending the checkpoint, src/server.py:92-97).  ``src``: the
        rank that holds the global model right now (the coordinator names it; default: this rank's view)."""
        src = self.primary_rank if src is None else int(src)
        if self.peers.buffers[src].base == 0:
            raise _lib.FedbError(f"the memory of rank {src} (holder of the global model) is not mapped here")
        # also into this rank's own aux arena: should the roles fall to this rank before the next committed round, it
        # already holds a valid global model for the next newcomer
        K.broadcast(self.peers.buffers[src].aux_arena(self.man), [self.trainer.arena, self.peers.me.aux_arena(self.man)])
        if self.codec == "fp8":
            # the pulled model IS the common base of the next round's deltas (the survivors hold the same one)
            self.base.copy_(self.trainer.arena.f32[:self.man.n_f32])
            self.residual.zero_()
            self._common_base = True
        torch.cuda.synchronize(self.device)

    def abort(self) -> None:
        """Host-side abort: breaks every spin on this rank's control block."""
        c = self.peers.me.ctrl()
        c[_lib.CTRL_OFF["abort_flag"] // 4] = 1

    def close(self):
        torch.cuda.synchronize(self.device)
        self.peers.close()
