"""Standalone trainer (fedb200/main.py): the single-node loop the reference keeps, unused, next to its federated one
(src/main.py:104-125, 193-228): train over the whole set, evaluate, keep the best checkpoint, cosine schedule, resume."""
import math

import torch

from fedb200 import main as M
from fedb200.utils import checkpoint as ckpt
from fedb200.utils import data as D


def test_cosine_matches_torch_scheduler():
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([p], lr=0.1)
    sch = torch.optim.lr_scheduler.CosineAnnealingLR(opt, T_max=200)
    for epoch in range(0, 60, 7):
        while sch.last_epoch < epoch:
            opt.step()
            sch.step()
        assert math.isclose(M.cosine_lr(0.1, epoch, 200), opt.param_groups[0]["lr"], rel_tol=1e-9, abs_tol=1e-12)


def test_epochs_best_checkpoint_and_resume(tmp_path):
    ds = D.synthetic_cifar(128 * 3, 100, seed=1)
    args = ["--model", "LeNet", "--epochs", "2", "--name", "solo", "--checkpoint-root", str(tmp_path), "--lr", "0.05",
            "--metrics", str(tmp_path / "m.jsonl")]
    out = M.run(args, dataset=ds)
    state = ckpt.load_checkpoint(out["checkpoint"])
    assert set(state) == {"net", "acc", "epoch"} and list(state["net"])[0] == "conv1.weight"      # reference layout
    assert math.isclose(state["acc"], out["best_acc"]) and state["epoch"] in (0, 1)
    assert out["last"]["steps"] == 3 and out["last"]["epoch"] == 1
    assert math.isclose(out["last"]["lr"], M.cosine_lr(0.05, 1, 200))
    # resume continues from the stored epoch and never lowers the best accuracy
    out2 = M.run(args + ["-r", "--no-schedule"], dataset=ds)
    assert out2["last"]["epoch"] == state["epoch"] + 1 and out2["best_acc"] >= out["best_acc"]
    assert out2["last"]["lr"] == 0.05
