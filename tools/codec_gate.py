#!/usr/bin/env python
"""Accuracy gate of the lossy `-c` codec (BASELINE north star (4)).

The reference's `-c Y` is lossless gzip on the gRPC channel (src/server.py:103-107, src/client.py:39-43).  On the NVLink
plane "lossless" is raw fp32 words; the opt-in `--codec fp8` ships block-scaled e4m3 deltas with error feedback instead.
A throughput number with fp8 only counts if the TRAINED model is as good as the lossless one, so this tool trains the
same federation twice - same seeds, same data, same number of rounds, `codec=none` vs `codec=fp8` - on a learnable
synthetic task and compares the evaluation curves:

    gate = over the last 5 rounds, |mean accuracy difference| <= 0.5 points  AND  mean eval loss within 2 %

plus the relative L2 distance of the final global models (informative).  Launch like bench.py:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \\
        tools/codec_gate.py --models LeNet ResNet18 --rounds 20 --out profiles/codec_gate.json
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--models", nargs="+", default=["LeNet", "ResNet18"])
    ap.add_argument("--rounds", type=int, default=20)
    ap.add_argument("--n-train", type=int, default=25600)
    ap.add_argument("--n-test", type=int, default=5000)
    ap.add_argument("--signal", type=float, default=0.1)
    ap.add_argument("--lr", type=float, default=0.05)
    ap.add_argument("--out", default="")
    a = ap.parse_args()

    import torch
    import torch.distributed as dist
    from fedb200.parallel.session import GpuFederation
    from fedb200.utils import data as D

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ds = D.synthetic_cifar(a.n_train, a.n_test, seed=0, learnable=True, signal=a.signal)
    report = {"clients": world, "rounds": a.rounds, "train_samples": a.n_train, "test_samples": a.n_test,
              "data": f"synthetic, class-conditional patterns, signal {a.signal}", "lr": a.lr, "models": {}}
    ok_all = True
    for model in a.models:
        curves, finals = {}, {}
        # "none2" repeats the lossless arm: split-K atomics make a zoo model's training non-reproducible bit for bit, so the
        # distance between two lossless runs is the noise floor the fp8 arm has to be read against
        for arm in ("none", "none2", "fp8"):
            codec = "fp8" if arm == "fp8" else "none"
            fed = GpuFederation(model=model, rank=rank, world=world, device=local, dataset=ds, seed=1000 + rank, lr=a.lr,
                                codec=codec, eval_mode="inline",
                                tag=f"gate-{model}-{arm}-{os.environ.get('MASTER_PORT', 'solo')}-{os.getppid()}")
            rows = []
            for r in range(a.rounds):
                tr, ev = fed.run_round(upload=(r == 0), sync=True)
                rows.append({"round": r + 1, "train_loss": tr.mean_loss, "eval_loss": ev.mean_loss, "eval_acc": ev.accuracy})
                if rank == 0:
                    print(f"[{model} {arm}] round {r + 1}: train {tr.mean_loss:.4f} eval {ev.mean_loss:.4f} "
                          f"acc {ev.accuracy:.4f}", flush=True)
            curves[arm] = rows
            n = fed.group.man.n_f32
            finals[arm] = fed.trainer.arena.f32[:n].detach().cpu().clone()
            fed.close()
            if world > 1:
                dist.barrier()
        tail = slice(-5, None)
        acc = {c: sum(r["eval_acc"] for r in curves[c][tail]) / 5 for c in curves}
        loss = {c: sum(r["eval_loss"] for r in curves[c][tail]) / 5 for c in curves}
        d_acc = abs(acc["fp8"] - acc["none"]) * 100.0
        d_loss = abs(loss["fp8"] - loss["none"]) / max(loss["none"], 1e-12)
        dist_rel = float((finals["fp8"] - finals["none"]).norm() / finals["none"].norm())
        noise = {"delta_acc_points": abs(acc["none2"] - acc["none"]) * 100.0,
                 "delta_loss_rel": abs(loss["none2"] - loss["none"]) / max(loss["none"], 1e-12),
                 "final_weight_rel_l2": float((finals["none2"] - finals["none"]).norm() / finals["none"].norm())}
        passed = d_acc <= 0.5 and d_loss <= 0.02
        ok_all &= passed
        report["models"][model] = {"curves": curves, "last5_acc": acc, "last5_eval_loss": loss,
                                   "delta_acc_points": d_acc, "delta_loss_rel": d_loss,
                                   "final_weight_rel_l2": dist_rel, "lossless_run_to_run": noise, "pass": passed}
        if rank == 0:
            print(f"[{model}] last-5 acc none {acc['none']:.4f} fp8 {acc['fp8']:.4f} (d {d_acc:.2f} pt), "
                  f"loss none {loss['none']:.4f} fp8 {loss['fp8']:.4f} (d {100 * d_loss:.2f} %), "
                  f"|w_fp8 - w_none| / |w_none| = {dist_rel:.3e} -> {'PASS' if passed else 'FAIL'}; two lossless runs differ by "
                  f"{noise['delta_acc_points']:.2f} pt / {100 * noise['delta_loss_rel']:.2f} % / {noise['final_weight_rel_l2']:.3e}",
                  flush=True)
    report["gate"] = "|d acc| <= 0.5 pt and eval loss within 2 % over the last 5 rounds"
    report["pass"] = bool(ok_all)
    if rank == 0:
        if a.out:
            with open(a.out, "w") as f:
                json.dump(report, f, indent=1)
        print(json.dumps({k: v for k, v in report.items() if k != "models"} |
                         {"summary": {m: {k: v for k, v in r.items() if k != "curves"} for m, r in report["models"].items()}}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
