"""``CpuFederation`` - the rank-symmetric round API of ``GpuFederation`` on hosts without a GPU.

Same contract (one process per client, ``run_round()`` = local pass over shard ``(rank, world)`` [N9] -> FedAvg in the
reference's order [N1-N3] over the LIVE ranks -> every live client evaluates the global model [N8]); the transport is
``torch.distributed`` with the ``gloo`` backend instead of NVLink peer memory: every rank all-gathers the flat arenas
(one fp32 buffer + one int64 buffer per client - the same two buffers K-AGG reads in place on the GPU plane) and
reduces them locally with the oracle ``fedavg_flat_cpu``, so all ranks hold bit-identical global models without a
server in the middle.  Reference: src/server.py:113-179 (round loop + allreduce), src/client.py:16-31.

It exists for three reasons: users of the reference run on CPU boxes (BASELINE config 1); the host-side logic of a
session (shard rule, live sets, sample-count weights, stats) is testable under ``gloo`` with world_size > 1 on a box
without GPUs; and it is the CPU oracle of a whole federated run for the GPU plane's tests.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.distributed as dist

from ..control.trainer import LocalTrainer, TrainerConfig, TrainStats
from ..utils import data as D
from .fedavg import fedavg_flat_cpu


class CpuFederation:
    def __init__(self, model: str = "LeNet", rank: Optional[int] = None, world: Optional[int] = None,
                 dataset: Optional[D.ImageDataset] = None, seed: Optional[int] = None, lr: float = 0.1, batch_size: int = 128,
                 augment: bool = False, evaluate: bool = True, weighted: bool = False, max_batches: Optional[int] = None,
                 codec: str = "none"):
        """``codec="fp8"``: the lossy, opt-in meaning of ``-c`` (block-scaled e4m3 deltas against the last global model, error
        feedback; ``fedb200/parallel/codec.py`` - the very arithmetic of K-CODEC on the GPU plane): one byte + 1/32 scale per
        weight crosses the wire instead of four bytes.  The first aggregation is lossless and establishes the common base
        (the workers start from different random models, N6); int64 counters always travel exactly."""
        if world is None:
            world = dist.get_world_size() if dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank() if dist.is_initialized() else 0
        if world > 1 and not dist.is_initialized():
            raise RuntimeError("CpuFederation with world > 1 needs torch.distributed (backend 'gloo') to be initialised")
        self.rank, self.world, self.evaluate, self.weighted = rank, world, evaluate, weighted
        ds = dataset if dataset is not None else D.synthetic_cifar(seed=0)
        cfg = TrainerConfig(model=model, lr=lr, batch_size=batch_size, device="cpu", seed=(rank if seed is None else seed),
                            augment=augment, max_batches=max_batches)
        self.trainer = LocalTrainer(cfg, ds)
        self.live: List[int] = list(range(world))
        self.rounds = 0
        if codec not in ("none", "fp8"):
            raise ValueError(f"codec {codec!r} (none | fp8)")
        self.codec = codec
        self._base = None                 # last global model: what the deltas are relative to (identical on every rank)
        self._residual = None             # what quantisation dropped so far on THIS rank (error feedback)
        self.bytes_sent_last = 0

    # ------------------------------------------------------------------ membership
    def set_live(self, live: Sequence[int]) -> None:
        """Only these ranks train, are averaged and evaluate in the next rounds (every rank must be told the same set)."""
        self.live = sorted(int(r) for r in live)

    # ------------------------------------------------------------------ aggregation
    def _gather(self, t: torch.Tensor) -> List[torch.Tensor]:
        if self.world == 1:
            return [t]
        out = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(out, t.contiguous())
        return out

    def aggregate(self, samples: int = 0) -> None:
        """FedAvg over the live ranks, result in every live rank's arena.  Dead ranks still take part in the collective
        (gloo needs every process) but their buffers are ignored and their arenas stay untouched."""
        a = self.trainer.arena
        man = a.manifest
        n = man.n_f32
        if self.codec == "fp8" and self._base is not None and not self.weighted:
            from . import codec as C
            q, s = C.encode(a.f32[:n], self._base, self._residual)
            qs, ss = self._gather(q), self._gather(s)
            i64s = self._gather(a.i64[:man.n_i64])
            self.bytes_sent_last = q.numel() + s.numel() * 4 + man.n_i64 * 8
            if self.rank in self.live:
                with torch.no_grad():
                    a.f32[:n].copy_(C.fedavg_decoded(self._base, [qs[r] for r in self.live], [ss[r] for r in self.live]))
                    i64 = i64s[self.live[0]].clone()
                    for r in self.live[1:]:
                        i64 += i64s[r]
                    a.i64[:man.n_i64].copy_((i64.to(torch.float32) / float(len(self.live))).trunc().to(torch.int64))
                self._base = a.f32[:n].detach().clone()
            return
        f32s = self._gather(a.f32[:n])
        i64s = self._gather(a.i64[:man.n_i64])
        self.bytes_sent_last = n * 4 + man.n_i64 * 8
        weights = None
        if self.weighted:
            counts = self._gather(torch.tensor([samples], dtype=torch.int64))
            weights = [int(counts[r][0]) for r in self.live]
        if self.rank in self.live:
            fedavg_flat_cpu([f32s[r] for r in self.live], [i64s[r] for r in self.live], weights,
                            a.f32[:n], a.i64[:man.n_i64])
            if self.codec == "fp8":
                self._base = a.f32[:n].detach().clone()
                self._residual = torch.zeros(n)

    # ------------------------------------------------------------------ one round
    def run_round(self):
        """Returns (train_stats, eval_stats) of this rank; ranks outside the live set return empty stats."""
        train, ev = TrainStats(), TrainStats()
        if self.rank in self.live:
            train = self.trainer.train(1, self.live.index(self.rank), self.world)      # dense rank among the live, world = total
        self.aggregate(train.samples)
        if self.evaluate and self.rank in self.live:
            ev = self.trainer.test()
        self.rounds += 1
        return train, ev

    def state_dict(self):
        return self.trainer.state_dict()

    def close(self):
        pass
