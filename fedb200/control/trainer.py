"""Local trainer - one worker's model, optimizer state, data and checkpoint.

Behaviour kept from the reference (src/main.py:30-242):
* SGD(lr=0.1, momentum=0.9, weight_decay=5e-4), PyTorch semantics (first step seeds the
  momentum buffer with the gradient, no dampening, no Nesterov); the LR is constant - the
  reference creates a cosine schedule and never steps it (N5).
* optimizer state is LOCAL and PERSISTENT: momentum survives rounds and is never averaged (N4).
* ``train(epoch, rank, world)``: one pass over the unshuffled batch sequence (batch 128),
  training only batches with ``(b+1) % world == rank`` (N9), CrossEntropy loss.
* ``test()``: the whole test set in eval mode, batch 100; returns loss/accuracy.
* every worker seeds nothing by default, so round 1 averages different random inits (N6); pass
  ``seed`` for reproducible runs.

Changed on purpose: no import-time side effects (B8), state lives in a flat arena instead of being
re-read from a checkpoint file before every pass, loss/accuracy are accumulated on the device and
read once per pass instead of two host syncs per step (K13), checkpoint directory is created on
demand (B7), the model is a flag (B10).
"""
from __future__ import annotations

import threading
import time
from dataclasses import dataclass, field
from typing import Optional

import torch

from .. import models
from ..ops import functional as F
from ..parallel.arena import Arena
from ..utils import checkpoint as ckpt
from ..utils import data as D
from ..utils.metrics import ProgressBar


@dataclass
class TrainStats:
    samples: int = 0
    steps: int = 0
    loss_sum: float = 0.0
    correct: int = 0
    seconds: float = 0.0

    @property
    def mean_loss(self):
        return self.loss_sum / max(self.steps, 1)

    @property
    def accuracy(self):
        return self.correct / max(self.samples, 1)


@dataclass
class TrainerConfig:
    model: str = "MobileNet"
    lr: float = 0.1
    momentum: float = 0.9
    weight_decay: float = 5e-4
    batch_size: int = 128
    test_batch_size: int = 100
    augment: bool = False           # reference augments; parity runs use deterministic data
    device: str = "cpu"
    seed: Optional[int] = None
    checkpoint_path: Optional[str] = None
    resume: bool = False
    use_cuda_graph: bool = True
    max_batches: Optional[int] = None   # truncate the pass (tests)
    progress: bool = False              # reference-style progress bar per batch on a TTY (src/main.py:157-158, utils.py:51-92)


def sgd_step_flat_(p: torch.Tensor, g: torch.Tensor, v: torch.Tensor, lr: float, momentum: float,
                   weight_decay: float, first: bool) -> None:
    """torch.optim.SGD on flat buffers: g += wd*p; v = g (first) | mu*v + g; p -= lr*v."""
    g = g.add(p, alpha=weight_decay) if weight_decay != 0 else g
    if momentum != 0:
        if first:
            v.copy_(g)
        else:
            v.mul_(momentum).add_(g)
        g = v
    p.add_(g, alpha=-lr)


class LocalTrainer:
    def __init__(self, cfg: TrainerConfig, dataset: Optional[D.ImageDataset] = None):
        self.cfg = cfg
        self.device = torch.device(cfg.device)
        if cfg.seed is not None:
            torch.manual_seed(cfg.seed)
        self.model = models.build(cfg.model).to(self.device)
        self.arena = Arena.adopt(self.model, self.device, with_optimizer_state=True)
        self.dataset = dataset if dataset is not None else D.synthetic_cifar(seed=0)
        self.lock = threading.Lock()        # one pass at a time on this model (fixes B2)
        self.best_acc, self.start_epoch = 0.0, 0
        self._aug_gen = torch.Generator().manual_seed(0 if cfg.seed is None else cfg.seed)
        self._engine = None
        if cfg.resume and cfg.checkpoint_path:
            state = ckpt.load_checkpoint(cfg.checkpoint_path)
            self.arena.load_state_dict(state["net"])
            self.best_acc, self.start_epoch = state["acc"], state["epoch"]
        if self.device.type == "cuda":
            from ..ops.engine import CudaEngine
            self._engine = CudaEngine(self)

    # ------------------------------------------------------------------ state movement
    def state_dict(self):
        return self.arena.state_dict(to_cpu=True)

    def load_state_dict(self, sd):
        self.arena.load_state_dict(sd)

    def save(self, path: Optional[str] = None, epoch: int = 1, acc=1):
        path = path or self.cfg.checkpoint_path
        if path:
            ckpt.save_checkpoint(path, self.state_dict(), acc=acc, epoch=epoch)

    def dumps(self, epoch: int = 1) -> bytes:
        return ckpt.dumps(self.state_dict(), acc=1, epoch=epoch)

    def set_lr(self, lr: float) -> None:
        """New learning rate for the following passes (standalone training with a schedule, fedb200/main.py).  The rate
        is an argument baked into captured graphs, so the GPU engine drops them and captures again on the next pass."""
        self.cfg.lr = float(lr)
        if self._engine is not None:
            self._engine.rebind_arena(self.arena)

    # ------------------------------------------------------------------ training
    def shard(self, rank: int, world: int):
        nb = self.dataset.num_train_batches(self.cfg.batch_size)
        if self.cfg.max_batches is not None:
            nb = min(nb, self.cfg.max_batches)
        return D.shard_batches(nb, rank, world)

    def _cpu_batch(self, b: int):
        bs = self.cfg.batch_size
        x = self.dataset.train_x[b * bs:(b + 1) * bs]
        y = self.dataset.train_y[b * bs:(b + 1) * bs]
        if self.cfg.augment:
            x = D.augment_batch_cpu(x, self._aug_gen)
        return D.normalize_batch(x), y

    def train(self, epoch: int, rank: int, world: int) -> TrainStats:
        with self.lock:
            t0 = time.perf_counter()
            batches = self.shard(rank, world)
            if self._engine is not None:
                st = self._engine.train_pass(batches)
            else:
                st = self._train_cpu(batches)
            st.seconds = time.perf_counter() - t0
            return st

    def _train_cpu(self, batches) -> TrainStats:
        cfg, a = self.cfg, self.arena
        st = TrainStats()
        self.model.train()
        bar = ProgressBar(len(batches)) if cfg.progress and batches else None
        for i, b in enumerate(batches):
            x, y = self._cpu_batch(b)
            a.zero_grad()
            out = self.model(x)
            loss = F.cross_entropy(out, y)
            loss.backward()
            with torch.no_grad():
                n = a.manifest.n_param
                sgd_step_flat_(a.f32[:n], a.grad, a.momentum, cfg.lr, cfg.momentum,
                               cfg.weight_decay, first=(a.sgd_steps == 0))
                a.sgd_steps += 1
                st.loss_sum += float(loss)
                st.correct += int((out.argmax(1) == y).sum())
            st.samples += y.numel()
            st.steps += 1
            if bar is not None:
                bar.update(i, "Loss: %.3f | Acc: %.3f%% (%d/%d)" % (st.mean_loss, 100.0 * st.accuracy, st.correct, st.samples))
        return st

    # ------------------------------------------------------------------ evaluation
    def test(self) -> TrainStats:
        with self.lock:
            t0 = time.perf_counter()
            if self._engine is not None:
                st = self._engine.eval_pass()
            else:
                st = self._test_cpu()
            st.seconds = time.perf_counter() - t0
            return st

    def _test_cpu(self) -> TrainStats:
        st = TrainStats()
        self.model.eval()
        bs = self.cfg.test_batch_size
        n = self.dataset.test_x.shape[0]
        bar = ProgressBar((n + bs - 1) // bs) if self.cfg.progress and n else None
        with torch.no_grad():
            for lo in range(0, n, bs):
                x = D.normalize_batch(self.dataset.test_x[lo:lo + bs])
                y = self.dataset.test_y[lo:lo + bs]
                out = self.model(x)
                st.loss_sum += float(F.cross_entropy(out, y))
                st.correct += int((out.argmax(1) == y).sum())
                st.samples += y.numel()
                st.steps += 1
                if bar is not None:
                    bar.update(lo // bs, "Loss: %.3f | Acc: %.3f%% (%d/%d)" % (st.mean_loss, 100.0 * st.accuracy, st.correct, st.samples))
        return st
