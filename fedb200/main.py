"""Standalone (non-federated) training of one zoo model: ``python -m fedb200.main --model LeNet --epochs 5``.

The reference's ``src/main.py`` carries, next to the federated ``train(e, rank, world)`` / ``test(e, count)`` that the
worker calls, the single-node loop it was derived from: ``train(epoch)`` over the whole train set, ``test(epoch)`` with
best-accuracy checkpointing to ``./checkpoint/<name>.pth`` and a ``CosineAnnealingLR(T_max=200)`` that is created but
never stepped (src/main.py:101-125, 193-228, 231-242).  This module is that loop on fedb200's trainer - hand-written
kernels and CUDA graphs on a GPU, stock PyTorch on the CPU - with the schedule actually stepped once per epoch
(``--no-schedule`` keeps the rate constant like the reference does in practice).  Flags of the reference's main:
``--lr`` and ``-r/--resume`` (src/main.py:20-28).
"""
from __future__ import annotations

import argparse
import math
import os
import sys
from typing import Optional, Sequence

from .control.trainer import LocalTrainer, TrainerConfig
from .utils import checkpoint as ckpt
from .utils import data as D
from .utils.metrics import JsonlSink


def parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="fedb200 standalone trainer (single node, no federation)")
    p.add_argument("--lr", type=float, default=0.1, help="learning rate (reference: src/main.py:21)")
    p.add_argument("-r", "--resume", action="store_true", help="resume from the checkpoint (reference: src/main.py:22-23)")
    p.add_argument("--model", default="MobileNet", help="zoo constructor (the reference hard-codes MobileNet, src/main.py:69)")
    p.add_argument("--epochs", type=int, default=1)
    p.add_argument("--name", default="ckpt", help="checkpoint stem: ./checkpoint/<name>.pth")
    p.add_argument("--checkpoint-root", default="./checkpoint")
    p.add_argument("--device", default="cpu")
    p.add_argument("--batch-size", type=int, default=128)
    p.add_argument("--t-max", type=int, default=200, help="CosineAnnealingLR period in epochs (reference: 200)")
    p.add_argument("--no-schedule", action="store_true", help="constant learning rate")
    p.add_argument("--augment", action="store_true", help="random crop + flip like the reference's train transform")
    p.add_argument("--dataset-size", type=int, default=50000, help="synthetic CIFAR-shaped train samples")
    p.add_argument("--test-size", type=int, default=10000)
    p.add_argument("--data-root", default=None, help="directory holding cifar-10-batches-py (real CIFAR-10)")
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--metrics", default=None, help="JSONL file for per-epoch records")
    p.add_argument("--progress", action="store_true", help="per-batch progress bar on a TTY (CPU trainer), like the reference")
    return p


def cosine_lr(base_lr: float, epoch: int, t_max: int) -> float:
    """torch.optim.lr_scheduler.CosineAnnealingLR with eta_min = 0, closed form."""
    return 0.5 * base_lr * (1.0 + math.cos(math.pi * epoch / t_max))


def run(argv: Optional[Sequence[str]] = None, dataset: Optional[D.ImageDataset] = None) -> dict:
    a = parser().parse_args(argv)
    path = ckpt.worker_checkpoint_path(a.name, a.checkpoint_root)
    ds = dataset if dataset is not None else D.get_dataset(a.data_root, a.dataset_size, a.test_size, seed=a.seed)
    cfg = TrainerConfig(model=a.model, lr=a.lr, batch_size=a.batch_size, augment=a.augment, device=a.device, seed=a.seed, progress=a.progress,
                        checkpoint_path=path, resume=a.resume and os.path.exists(path))
    t = LocalTrainer(cfg, ds)
    sink = JsonlSink(a.metrics)
    best, start = (float(t.best_acc), int(t.start_epoch)) if cfg.resume else (0.0, 0)
    last = None
    for epoch in range(start, start + a.epochs):
        if not a.no_schedule:
            t.set_lr(cosine_lr(a.lr, epoch, a.t_max))
        tr = t.train(epoch, 0, 1)                        # world of one: every batch is this trainer's (N9 with world = 1)
        ev = t.test()
        acc = 100.0 * ev.accuracy
        saved = acc > best
        if saved:                                        # best-accuracy checkpoint, reference layout {'net','acc','epoch'}
            t.save(path, epoch=epoch, acc=acc)
            best = acc
        last = sink.write("epoch", epoch=epoch, lr=t.cfg.lr, train_loss=tr.mean_loss, train_acc=tr.accuracy, steps=tr.steps,
                          test_loss=ev.mean_loss, test_acc=ev.accuracy, saved=saved, seconds=tr.seconds + ev.seconds)
        print(f"epoch {epoch}: lr {t.cfg.lr:.5f}  train loss {tr.mean_loss:.4f} acc {100 * tr.accuracy:.2f}%  "
              f"test loss {ev.mean_loss:.4f} acc {acc:.2f}%{'  (saved)' if saved else ''}", flush=True)
    return {"best_acc": best, "epochs": a.epochs, "checkpoint": path, "last": last}


def main() -> int:
    run(sys.argv[1:])
    return 0


if __name__ == "__main__":
    sys.exit(main())
