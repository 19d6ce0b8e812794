"""GPU step engine behind ``LocalTrainer`` (SURVEY C12 on the device).

Two paths, both hand-written kernels end to end:
* LeNet - the fused two-kernel step of csrc/lenet.cu.  A whole local pass (every step of the
  shard) is captured ONCE into a CUDA graph per (shard shape) and replayed each round: one graph
  launch per round, no host syncs inside (K13), uint8 images normalised on the fly (K14).
* any other zoo model - the model's own ``forward`` (which dispatches every op to
  ``fedb200.ops.cuda_nn``) under autograd, CE + fused SGD kernels, the step captured in a CUDA
  graph as well.

Per pass the shard's images/labels are copied host->device from pinned memory IN CHUNKS on a forked
copy stream; step i only waits for the chunk that holds its batch, so the copy engine runs under the
step kernels.  The fork/join is captured with the steps: the H2D copies are memcpy nodes of the round
graph (one graph launch = upload + every step of the pass).  Loss / accuracy counters are read back once.
"""
from __future__ import annotations

import contextlib
import gc
from typing import Dict, List, Optional, Tuple

import torch

from . import _lib
from . import cuda as K


@contextlib.contextmanager
def _quiet_gc():
    """No cyclic garbage collection while a stream capture is open: the collector may finalise objects of an EARLIER
    engine (captured graphs with their private pools, pinned buffers), and the CUDA calls of those finalisers
    (cudaFree, cudaFreeHost, cudaGraphExecDestroy) are illegal during a global-mode capture and invalidate it."""
    gc.collect()
    was = gc.isenabled()
    gc.disable()
    try:
        yield
    finally:
        if was:
            gc.enable()


class CudaEngine:
    def __init__(self, trainer):
        from ..control.trainer import TrainStats
        self.TrainStats = TrainStats
        self.t = trainer
        self.dev = trainer.device
        self.cfg = trainer.cfg
        self.arena = trainer.arena
        self.is_lenet = trainer.cfg.model == "LeNet"
        self.copy_stream = torch.cuda.Stream(device=self.dev)
        self._shards: Dict[Tuple[int, ...], dict] = {}
        self._test_dev = None
        self._test_resident = False
        self._tex = None                  # evaluation-set exchange over NVLink (multi-GPU sessions)
        self._slot = 0                    # half of the exchange double buffer that holds / receives the current round's set
        self._tex_on = False
        self._eval_graph = None
        self.snap = None                  # snapshot arena scored by the off-critical-path evaluation
        self.eval_model = None
        self.eval_stream = None
        self.h2d_bytes_last = 0
        self.d2h_bytes_last = 0
        self.seed_dev = torch.zeros(1, dtype=torch.int64, device=self.dev)
        self.eval_stats = torch.zeros(4, dtype=torch.float32, device=self.dev)
        self._pinned_stats = torch.zeros(8, dtype=torch.float32).pin_memory()
        self.rebind_arena(self.arena)

    def rebind_arena(self, arena) -> None:
        """(Re)attach to an arena - called again when the trainer's state moves into peer memory.
        Captured graphs hold raw pointers, so they are dropped."""
        self.arena = arena
        for sh in self._shards.values():
            sh["graph"] = None
        self._eval_graph = None
        # per-model caches of the convolution glue (packed filters, packed gradient accumulators): they hold addresses of
        # the arena, so they start over with it
        from . import cuda_gemm
        self._grad_clean = False
        self.scope = cuda_gemm.StepScope(self.dev, train=True)
        self.eval_scope = cuda_gemm.StepScope(self.dev, train=False)
        with torch.cuda.device(self.dev):
            if self.is_lenet:
                self.plan = K.LeNetPlan(arena, self.cfg.lr, self.cfg.momentum, self.cfg.weight_decay,
                                        augment=self.cfg.augment, max_batch=self.cfg.batch_size)
                self.plan.desc.seed_dev = self.seed_dev.data_ptr()
                self.stats = self.plan.stats
            else:
                self.stats = torch.zeros(4, dtype=torch.float32, device=self.dev)
            if self.snap is not None and self.is_lenet:
                self.plan.bind_eval_params(self.snap.f32)

    # ------------------------------------------------------------------ evaluation off the critical path
    def enable_snapshot_eval(self) -> None:
        """Evaluation of a round's global model from a private copy on a second stream, so that the next local pass
        (which overwrites the arena) does not have to wait for it.  The reference evaluates inside ``SendModel``
        (src/client.py:30, src/main.py:167-191), i.e. on the round's critical path; the same 10,000 images are still
        scored on every client every round here, only concurrently with the next round's training."""
        if self.snap is not None:
            return
        from .. import models
        from ..parallel.arena import Arena
        with torch.cuda.device(self.dev):
            if self.is_lenet:
                self.snap = Arena(self.arena.manifest, self.dev)
                self.plan.bind_eval_params(self.snap.f32)
            else:
                from . import cuda_gemm
                self.eval_model = models.build(self.cfg.model).to(self.dev)
                self.snap = Arena.adopt(self.eval_model, self.dev)
                self.eval_model.eval()
                self.snap_scope = cuda_gemm.StepScope(self.dev, train=False)
            self.eval_stream = torch.cuda.Stream(device=self.dev)
            self.eval_copy_stream = torch.cuda.Stream(device=self.dev)
            self.eval_stats_prev = torch.zeros(4, dtype=torch.float32, device=self.dev)
        self._eval_graph = None

    # ---- LeNet: the overlapped evaluation as a work queue shared by a background launch and a finisher
    def queue_eval_available(self, upload: bool) -> bool:
        """The fused LeNet kernels and a test set that is already on the device (with ``upload`` it rode in the training
        pass's graph; without, the resident copy of an earlier round is scored, as ``eval_pass`` does)."""
        return self.is_lenet and self.snap is not None and (self._test_resident or not upload) and self._test_dev is not None

    def queue_eval_start(self) -> None:
        """On the evaluation stream, after the snapshot: zero the chunk counter and start the background CTAs - as many
        as the training step leaves SMs idle (its sample kernel runs one CTA per sample of a batch, 128 of 148 SMs), so
        they never take an SM from it."""
        if not hasattr(self, "_q_counter"):
            self._q_counter = torch.zeros(1, dtype=torch.int32, device=self.dev)
            props = torch.cuda.get_device_properties(self.dev)
            self._sms = props.multi_processor_count
        imgs, labels = self._test_images(), self._test_dev[1]
        self._q = (imgs, labels, int(imgs.shape[0]))
        self._test_resident = False
        K.zero_(self._q_counter)
        import os
        idle = max(self._sms - self.cfg.batch_size, 0)
        idle = int(os.environ.get("FEDB200_EVAL_BG_CTAS", idle))
        if idle > 0:
            self.plan.eval_queue(imgs, labels, self._q_counter, idle, snapshot=True)

    def queue_eval_finish(self) -> None:
        """On the training stream, when the result is needed: a full-size launch takes the chunks that are left (none
        when the background CTAs had a whole local pass to themselves)."""
        imgs, labels, n = self._q
        self.plan.eval_queue(imgs, labels, self._q_counter, 2 * self._sms, snapshot=True)

    def join_overlapped_eval(self, keep: bool = False) -> None:
        """The current stream waits for the evaluation in flight on the second stream (a queue evaluation is first helped
        to its end by a full-size launch, then reduced).  In steady state it ended long ago (it is a fraction of a local
        pass); the wait is what makes it safe to overwrite the snapshot and, two rounds later, the half of the test-set
        double buffer it read.  ``keep`` parks the finished counters in ``eval_stats_prev`` for the round's result copy.
        Idempotent."""
        ev = getattr(self, "_eval_done", None)
        if ev is None:
            return
        cur = torch.cuda.current_stream(self.dev)
        queued = getattr(self, "_q", None) is not None
        if queued:
            self.queue_eval_finish()         # takes what the background CTAs have not scored yet, at full width
        cur.wait_event(ev)
        if queued:
            self.queue_eval_stats()
        if keep:
            self.eval_stats_prev.copy_(self.eval_stats, non_blocking=True)

    def queue_eval_stats(self) -> None:
        """After BOTH launches have ended: fixed-order reduction of the per-sample rows into ``eval_stats``."""
        K.zero_(self.eval_stats)
        self.plan.eval_queue_stats(self._q[2], self.eval_stats)
        self._q = None

    def snapshot(self) -> None:
        """arena -> snapshot on the current stream (one K-BCAST launch over local memory)."""
        K.broadcast(self.arena, [self.snap])

    # ------------------------------------------------------------------ data staging
    def _shard(self, batches: List[int]) -> dict:
        key = tuple(batches)
        sh = self._shards.get(key)
        if sh is None:
            bs = self.cfg.batch_size
            ds = self.t.dataset
            idx = torch.cat([torch.arange(b * bs, min((b + 1) * bs, ds.train_x.shape[0])) for b in batches]) \
                if batches else torch.zeros(0, dtype=torch.long)
            hx = ds.train_x[idx].contiguous()
            hy = ds.train_y[idx].contiguous()
            if torch.cuda.is_available():
                hx, hy = hx.pin_memory(), hy.pin_memory()
            sizes = [min((b + 1) * bs, ds.train_x.shape[0]) - b * bs for b in batches]
            sh = {"hx": hx, "hy": hy, "sizes": sizes,
                  "dx": torch.empty_like(hx, device=self.dev), "dy": torch.empty_like(hy, device=self.dev),
                  "graph": None}
            if len(self._shards) > 8:
                self._shards.clear()
            self._shards[key] = sh
        return sh

    def _h2d(self, dst: torch.Tensor, src: torch.Tensor, lo: int, hi: int, stream) -> None:
        """rows [lo, hi) of a pinned host tensor -> the same rows of its device twin, as ONE async copy."""
        if hi <= lo:
            return
        row = src[0].numel() * src.element_size() if src.dim() > 1 else src.element_size()
        _lib.check(_lib.lib().fedb_copy_async(dst.data_ptr() + lo * row, src.data_ptr() + lo * row, (hi - lo) * row,
                                              stream.cuda_stream), "fedb_copy_async")

    def _streamed(self, pairs, bounds: List[int], work, upload: bool, tail=None, side=None) -> None:
        """Runs work(i) for every chunk i = rows [bounds[i], bounds[i+1]) on the current stream.  With ``upload`` the
        rows of every (device, pinned) pair are first copied H2D chunk by chunk on the copy stream, and work(i) waits
        for chunk i only; ``tail()`` then queues further copies behind them (they finish under the remaining work).
        Works live and under stream capture (fork / join through events)."""
        cur = torch.cuda.current_stream(self.dev)
        if not upload:
            for i in range(len(bounds) - 1):
                work(i)
            return
        side = self.copy_stream if side is None else side
        side.wait_stream(cur)                                         # fork
        ready = []
        for i in range(len(bounds) - 1):
            for d, h in pairs:
                self._h2d(d, h, bounds[i], bounds[i + 1], side)
            ev = torch.cuda.Event()
            ev.record(side)
            ready.append(ev)
        if tail is not None:
            tail(side)
        for i, ev in enumerate(ready):
            cur.wait_event(ev)
            work(i)
        cur.wait_stream(side)                                         # join

    def _test_buffers(self):
        if self._test_dev is None:
            ds = self.t.dataset
            hx, hy = ds.test_x.contiguous(), ds.test_y.contiguous()
            if torch.cuda.is_available():
                hx, hy = hx.pin_memory(), hy.pin_memory()
            self._test_host = (hx, hy)
            self._test_dev = (torch.empty_like(hx, device=self.dev), torch.empty_like(hy, device=self.dev))
        if self._tex is not None:          # images live in the peer-mapped double buffer, labels stay local
            return [(self._tex.images(self._slot), self._test_host[0]), (self._test_dev[1], self._test_host[1])]
        return list(zip(self._test_dev, self._test_host))

    def set_test_exchange(self, ex) -> None:
        """``ex`` (fedb200.parallel.session.TestExchange): the evaluation images are all-gathered over NVLink - every rank
        uploads 1/N of them and pushes that slice into every peer's double buffer - instead of N full PCIe uploads."""
        self._tex = ex
        self._eval_graph = None
        for sh in self._shards.values():
            sh["graph"] = None

    def _test_images(self) -> torch.Tensor:
        self._test_buffers()
        return self._tex.images(self._slot) if self._tex is not None else self._test_dev[0]

    def _test_bytes(self) -> int:
        return self._test_host[0].numel() + self._test_host[1].numel() * 8

    # ------------------------------------------------------------------ training pass
    COPY_CHUNK_STEPS = 8        # batches per H2D chunk in steady state (LeNet: 3 MB); the first chunks are 1, 2 and 4
                                # batches so that step 0 waits for one batch only, not for a whole chunk

    def _issue_steps(self, sh: dict, upload: bool = False, with_test: bool = False) -> None:
        sizes = sh["sizes"]
        offs = [0]
        for n in sizes:
            offs.append(offs[-1] + n)
        first, i, g = [], 0, 1                                       # first step of every chunk: 0, 1, 3, 7, 15, 23, ...
        while i < len(sizes):
            first.append(i)
            i += g
            g = min(2 * g, self.COPY_CHUNK_STEPS)
        bounds = [offs[f] for f in first] + [offs[-1]]
        last = first[1:] + [len(sizes)]

        def work(c):
            for i in range(first[c], last[c]):
                x, y = sh["dx"][offs[i]:offs[i + 1]], sh["dy"][offs[i]:offs[i + 1]]
                if self.is_lenet:
                    self.plan.train_step(x, y, step_seed=i + 1)
                else:
                    self._generic_step(x, y, i + 1)

        def tail(side):                                              # the round's test set rides behind the shard
            pairs = self._test_buffers()
            if self._tex is not None and self._tex_on:
                # all-gather: my 1/N of the images comes up from my host, K-PUSH sends it to every peer over NVLink
                # (their slices arrive the same way; this round's K-AGG flags order them before anybody's evaluation)
                lo, hi = self._tex.rows(self._tex.rank)
                self._h2d(pairs[0][0], pairs[0][1], lo, hi, side)
                self._h2d(pairs[1][0], pairs[1][1], 0, pairs[1][1].shape[0], side)
                self._tex.push(self._slot, side.cuda_stream)
            else:
                for d, h in pairs:
                    self._h2d(d, h, 0, h.shape[0], side)

        self._streamed([(sh["dx"], sh["hx"]), (sh["dy"], sh["hy"])], bounds, work, upload,
                       tail if (upload and with_test) else None)

    def _generic_step(self, x_u8, y, seed):
        a, cfg = self.arena, self.cfg
        with self.scope as sc:
            sc.fused_counters = True
            sc.begin()                                   # every packed filter of the model, one launch
            x = K.prepare_batch(x_u8, nhwc=True, augment=cfg.augment, seed=seed, seed_dev=self.seed_dev)
            if not self._grad_clean:                     # afterwards every step's SGD kernel leaves the gradient zeroed
                K.zero_(a.grad)
                self._grad_clean = True
            out = self.t.model(x)
            dlogits = torch.empty_like(out)
            K.cross_entropy_raw(out.detach().contiguous(), y, dlogits, None, self.stats)
            out.backward(dlogits)
            sc.end_backward()                            # every packed weight gradient into the arena, one launch
        n = a.manifest.n_param
        # the int64 segment holds the BatchNorm step counters, densely from 0 (its tail is alignment padding)
        n_ctr = sum(e.numel for e in a.manifest.entries if e.segment == "i64")
        K.sgd_step_end(a.f32[:n], a.grad, a.momentum, cfg.lr, cfg.momentum, cfg.weight_decay, counters=a.i64[:n_ctr])

    def train_pass(self, batches: List[int], upload: bool = True, sync: bool = True, with_test: bool = False):
        """One local pass over the shard.  ``upload``: this pass's inputs are copied from pinned memory inside the
        pass (chunked, overlapped); ``with_test``: the test set of the evaluation that follows the aggregation is
        copied in the same pass, behind the shard, so that the next ``eval_pass(upload=True)`` finds it resident."""
        with torch.cuda.device(self.dev):
            self.t.model.train()
            sh = self._shard(batches)
            if not sh["sizes"]:
                return self.TrainStats()
            upload = bool(upload)
            with_test = bool(with_test) and upload
            if with_test and self._tex is not None:
                self._slot ^= 1                               # this round's set goes into the other half
                self._tex_on = self._tex.enabled()
            key = (upload, with_test, self._slot if with_test else -1, self._tex_on if with_test else False)
            K.zero_(self.stats)
            self.seed_dev += 1
            if self.cfg.use_cuda_graph:
                graphs = sh["graph"] if sh["graph"] is not None else {}
                sh["graph"] = graphs
                if key not in graphs:
                    self._issue_steps(sh, upload, with_test)   # warm-up (also sizes every workspace)
                    self.scope.sync_tables()
                    torch.cuda.synchronize(self.dev)
                    g = torch.cuda.CUDAGraph()
                    # the warm-up already trained one pass; that is a real pass, so count it and
                    # capture the graph (with or without the H2D branch) for the following rounds
                    with _quiet_gc(), torch.cuda.graph(g, stream=torch.cuda.Stream(device=self.dev)):
                        self._issue_steps(sh, upload, with_test)
                    graphs[key] = g
                else:
                    graphs[key].replay()
                    _lib.count_launch(self._launches_per_pass(sh))
            else:
                self._issue_steps(sh, upload, with_test)
            if upload:
                self.h2d_bytes_last = sh["hx"].numel() + sh["hy"].numel() * 8
                if with_test:
                    self.h2d_bytes_last += self._test_bytes() if not (self._tex is not None and self._tex_on) \
                        else self._tex.slice_bytes() + self._test_host[1].numel() * 8
            self._test_resident = with_test
            self.arena.sgd_steps += len(sh["sizes"])
            return self.read_stats() if sync else None

    def _launches_per_pass(self, sh) -> int:
        return (2 if self.is_lenet else 0) * len(sh["sizes"])

    def read_stats(self, which: str = "train"):
        """One D2H read per pass: loss, correct, samples, steps."""
        s = (self.stats if which == "train" else self.eval_stats).cpu()
        self.d2h_bytes_last = s.numel() * 4
        return self._to_stats(s)

    def _to_stats(self, s):
        st = self.TrainStats()
        st.loss_sum, st.correct, st.samples, st.steps = float(s[0]), int(s[1]), int(s[2]), int(s[3])
        return st

    # ---- pipelined result read-back: the copy is enqueued with the round, the host reads it one round later
    def enqueue_round_stats(self, status_word: Optional[torch.Tensor] = None, eval_src: Optional[torch.Tensor] = None):
        """Queues the D2H copy of this round's counters (train 4 + eval 4 floats, plus an optional int32 status word
        such as the K-AGG flag status) into a pinned slot and returns a ticket for ``fetch_round_stats``.  Nothing
        blocks here, so the next round can be enqueued while this one still runs."""
        if not hasattr(self, "_ring"):
            self._ring = [(torch.zeros(8, dtype=torch.float32).pin_memory(), torch.zeros(1, dtype=torch.int32).pin_memory(),
                           torch.cuda.Event()) for _ in range(4)]
            self._ring_next = 0
        buf, st, ev = self._ring[self._ring_next]
        self._ring_next = (self._ring_next + 1) % len(self._ring)
        with torch.cuda.device(self.dev):
            buf[:4].copy_(self.stats, non_blocking=True)
            buf[4:].copy_(self.eval_stats if eval_src is None else eval_src, non_blocking=True)
            if status_word is not None:
                st.copy_(status_word, non_blocking=True)
            else:
                st.zero_()
            ev.record(torch.cuda.current_stream(self.dev))
        self.d2h_bytes_last = 36
        return buf, st, ev

    def fetch_round_stats(self, ticket):
        buf, st, ev = ticket
        ev.synchronize()
        return self._to_stats(buf[:4]), self._to_stats(buf[4:]), int(st[0]) & 0xFFFFFFFF

    def read_round_stats(self):
        """Train + eval counters of a round in ONE pinned D2H copy (32 bytes) and one sync."""
        self.join_overlapped_eval()
        self._pinned_stats[:4].copy_(self.stats, non_blocking=True)
        self._pinned_stats[4:].copy_(self.eval_stats, non_blocking=True)
        torch.cuda.current_stream(self.dev).synchronize()
        self.d2h_bytes_last = 32
        return self._to_stats(self._pinned_stats[:4]), self._to_stats(self._pinned_stats[4:])

    # ------------------------------------------------------------------ evaluation pass
    def eval_pass(self, upload: bool = True, sync: bool = True, chunk: int = 2000, snapshot: bool = False):
        """Evaluation of the current model on the whole test set.  ``upload`` copies the test set H2D first, in chunks
        that overlap the evaluation kernels - unless the preceding ``train_pass(with_test=True)`` of this round already
        brought it over, in which case the pass is a single launch over the resident copy."""
        with torch.cuda.device(self.dev):
            model = self.eval_model if (snapshot and self.eval_model is not None) else self.t.model
            model.eval()
            ds = self.t.dataset
            pairs = self._test_buffers()
            side = self.eval_copy_stream if snapshot else None
            copy = bool(upload) and not self._test_resident
            self._test_resident = False
            if copy:
                self.h2d_bytes_last += self._test_bytes()
            K.zero_(self.eval_stats)
            n = ds.test_x.shape[0]
            if self.is_lenet:
                step = chunk if copy else n
                bounds = list(range(0, n, step)) + [n]

                imgs = self._test_images()

                def work(c):
                    lo, hi = bounds[c], bounds[c + 1]
                    self.plan.eval_batch(imgs[lo:hi], self._test_dev[1][lo:hi], self.eval_stats, snapshot=snapshot)

                graphs = self._eval_graph if self._eval_graph is not None else {}
                self._eval_graph = graphs
                gkey = (copy, self._slot, snapshot)
                if self.cfg.use_cuda_graph and gkey in graphs:
                    graphs[gkey].replay()
                    _lib.count_launch(2 * (len(bounds) - 1))
                else:
                    self._streamed(pairs, bounds, work, copy, side=side)
                    if self.cfg.use_cuda_graph:
                        torch.cuda.synchronize(self.dev)
                        g = torch.cuda.CUDAGraph()
                        with _quiet_gc(), torch.cuda.graph(g, stream=torch.cuda.Stream(device=self.dev)):
                            self._streamed(pairs, bounds, work, copy, side=side)
                        graphs[gkey] = g
            else:
                # eval-mode results do not depend on the batch size (BN uses running statistics): evaluate in batches of
                # at least 500 instead of the reference's 100 (src/main.py:54-55) and replay the pass as one CUDA graph
                bs = max(self.cfg.test_batch_size, 500)
                bounds = list(range(0, n, bs)) + [n]

                imgs = self._test_images()

                def work(c):
                    lo, hi = bounds[c], bounds[c + 1]
                    x = K.prepare_batch(imgs[lo:hi], nhwc=True)
                    out = model(x)
                    K.cross_entropy_raw(out.contiguous(), self._test_dev[1][lo:hi], None, None, self.eval_stats)

                graphs = self._eval_graph if self._eval_graph is not None else {}
                self._eval_graph = graphs
                gkey = (copy, self._slot, snapshot)          # graph key: which half of the double buffer is read
                esc = self.snap_scope if (snapshot and self.eval_model is not None) else self.eval_scope

                def run_pass():
                    esc.begin()                              # the weights changed since the last pass: repack once,
                    self._streamed(pairs, bounds, work, copy, side=side)     # not once per batch and layer

                with torch.no_grad(), esc:
                    if self.cfg.use_cuda_graph and gkey in graphs:
                        graphs[gkey].replay()
                    else:
                        run_pass()
                        if self.cfg.use_cuda_graph:
                            esc.sync_tables()
                            torch.cuda.synchronize(self.dev)
                            g = torch.cuda.CUDAGraph()
                            with _quiet_gc(), torch.cuda.graph(g, stream=torch.cuda.Stream(device=self.dev)):
                                run_pass()
                            graphs[gkey] = g
            return self.read_stats("eval") if sync else None
