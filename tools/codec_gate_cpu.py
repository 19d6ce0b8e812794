#!/usr/bin/env python
"""The accuracy gate of the fp8 codec WITHOUT run-to-run noise: the CPU federation (gloo, one process per client, one thread
each) is bit-reproducible, so two lossless runs are identical and any difference between the arms is the codec's.

Same arithmetic as the GPU plane: `fedb200/parallel/codec.py` is the bit-exact oracle of the K-CODEC kernels
(tests/test_gpu_kernels.py::test_codec_kernels_match_the_cpu_oracle_bit_for_bit).  LeNet only - a zoo model takes hours on CPUs.

    python tools/codec_gate_cpu.py --clients 4 --rounds 30 --out profiles/codec_gate_cpu.json
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _worker(rank, world, port, a, arm, out_dir):
    import torch
    import torch.distributed as dist
    from fedb200.parallel.cpu_session import CpuFederation
    from fedb200.utils import data as D
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ds = D.synthetic_cifar(a.n_train, a.n_test, seed=0, learnable=True, signal=a.signal)
    fed = CpuFederation(model=a.model, dataset=ds, seed=1000 + rank, lr=a.lr, codec="fp8" if arm == "fp8" else "none")
    if arm == "perturbed":
        # the lossless run again with every initial weight scaled by (1 + 1e-6): how far a perturbation of one part in a
        # million carries the metrics after the same number of rounds - the yardstick for the codec's difference
        with torch.no_grad():
            fed.trainer.arena.f32.mul_(1.0 + 1e-6)
    rows = []
    for r in range(a.rounds):
        tr, ev = fed.run_round()
        rows.append({"round": r + 1, "train_loss": tr.mean_loss, "eval_loss": ev.mean_loss, "eval_acc": ev.accuracy,
                     "bytes_sent": fed.bytes_sent_last})
        if rank == 0:
            print(f"[{arm}] round {r + 1}: train {tr.mean_loss:.4f} eval {ev.mean_loss:.4f} acc {ev.accuracy:.4f}", flush=True)
    if rank == 0:
        n = fed.trainer.arena.manifest.n_f32
        torch.save({"rows": rows, "final": fed.trainer.arena.f32[:n].clone()}, os.path.join(out_dir, f"{arm}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="LeNet")
    ap.add_argument("--clients", type=int, default=4)
    ap.add_argument("--rounds", type=int, default=30)
    ap.add_argument("--n-train", type=int, default=10240)
    ap.add_argument("--n-test", type=int, default=2000)
    ap.add_argument("--signal", type=float, default=0.1)
    ap.add_argument("--lr", type=float, default=0.05)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    import torch
    import torch.multiprocessing as mp
    work = tempfile.mkdtemp(prefix="fedb200-gate-")
    res = {}
    for arm in ("none", "none2", "perturbed", "fp8"):
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            port = s.getsockname()[1]
        mp.start_processes(_worker, args=(a.clients, port, a, arm, work), nprocs=a.clients, join=True, start_method="spawn")
        res[arm] = torch.load(os.path.join(work, f"{arm}.pt"), weights_only=False)
    tail = slice(-5, None)
    acc = {k: sum(r["eval_acc"] for r in v["rows"][tail]) / 5 for k, v in res.items()}
    loss = {k: sum(r["eval_loss"] for r in v["rows"][tail]) / 5 for k, v in res.items()}
    d_acc = abs(acc["fp8"] - acc["none"]) * 100
    d_loss = abs(loss["fp8"] - loss["none"]) / max(loss["none"], 1e-12)
    rel = lambda x, y: float((res[x]["final"] - res[y]["final"]).norm() / res[y]["final"].norm())   # noqa: E731
    report = {"plane": "cpu/gloo (bit-reproducible)", "model": a.model, "clients": a.clients, "rounds": a.rounds,
              "train_samples": a.n_train, "test_samples": a.n_test, "lr": a.lr,
              "data": f"synthetic, class-conditional patterns, signal {a.signal}",
              "gate": "|d acc| <= 0.5 pt and eval loss within 2 % over the last 5 rounds",
              "last5_acc": acc, "last5_eval_loss": loss, "delta_acc_points": d_acc, "delta_loss_rel": d_loss,
              "final_weight_rel_l2": rel("fp8", "none"),
              "lossless_run_to_run": {"delta_acc_points": abs(acc["none2"] - acc["none"]) * 100,
                                      "delta_loss_rel": abs(loss["none2"] - loss["none"]) / max(loss["none"], 1e-12),
                                      "final_weight_rel_l2": rel("none2", "none")},
              "lossless_perturbed_1e-6": {"delta_acc_points": abs(acc["perturbed"] - acc["none"]) * 100,
                                          "delta_loss_rel": abs(loss["perturbed"] - loss["none"]) / max(loss["none"], 1e-12),
                                          "final_weight_rel_l2": rel("perturbed", "none")},
              "bytes_per_round": {k: v["rows"][-1]["bytes_sent"] for k, v in res.items()},
              "pass": bool(d_acc <= 0.5 and d_loss <= 0.02),
              "curves": {k: v["rows"] for k, v in res.items()}}
    print(json.dumps({k: v for k, v in report.items() if k != "curves"}))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(report, f, indent=1)


if __name__ == "__main__":
    main()
