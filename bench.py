#!/usr/bin/env python
"""Headline benchmark: federated rounds per second (BASELINE.json, config 2).

One "step" = one whole federated round of the reference (src/server.py:120-153):
every client runs one local pass over its 1/N shard of a CIFAR-shaped synthetic train set
(50,000 images, batch 128, SGD momentum 0.9 / wd 5e-4), the clients' models are FedAvg-ed, the global
model goes to the backup and back to every client, and every client evaluates it on the 10,000-image
test set.  Model: LeNet, the zoo's small CNN.  N clients = N GPUs, one process per GPU.

  python bench.py --gpus 1 --steps 20 --warmup 3                       # ours, 1 GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \\
         --master-port 29500 bench.py --gpus 8 --steps 20 --warmup 3   # ours, 8 GPUs
  python bench.py --impl reference --gpus 2 --steps 3 --warmup 1       # unmodified reference

Prints ONE JSON line (rank 0).  `value` is device-timed (CUDA events, max over ranks) with the inputs
resident on the GPU; `e2e` is the same metric through ``GpuFederation.run_round_async()`` (wall clock) including the
H2D copy of every round's inputs from pinned memory and the D2H read of every round's loss/accuracy/status; the
host reads round r's result after it has enqueued round r+1.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "federated_rounds_per_sec"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    ap.add_argument("--model", default="LeNet")
    ap.add_argument("--n-train", type=int, default=50000)
    ap.add_argument("--n-test", type=int, default=10000)
    ap.add_argument("--batch-size", type=int, default=128)
    ap.add_argument("--codec", choices=("none", "fp8"), default="none")
    ap.add_argument("--nvls-reduce", action="store_true")
    ap.add_argument("--no-multicast", action="store_true")
    ap.add_argument("--repl", choices=("overlap", "fused"), default="overlap")
    ap.add_argument("--ref-gpu-workers", action="store_true", help="reference arm: R-gpu variant (DP shim)")
    ap.add_argument("--eval", choices=("overlap", "inline"), default="inline",
                    help="inline: every client scores the whole test set every round on the round's critical path, as the "
                         "reference does; overlap: on a second stream, concurrently with the next local pass (measured "
                         "slower for LeNet on B200, see GpuFederation)")
    ap.add_argument("--no-augment", action="store_true", help="skip RandomCrop+Flip (the reference arm always applies them)")
    ap.add_argument("--no-l2-flush", action="store_true", help="diagnostic: skip the L2 flush between device-timed rounds")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the timed region runs (B200_PROFILING recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = sorted({n for r in self.rows if len(r) >= 8 for n, v in zip(names, r[4:8]) if v.startswith("Active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def run_reference_arm(a) -> dict:
    from baseline.ref_run import reference_available, run_reference
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return {}
    why = reference_available()
    if why:
        return {"impl": "reference", "unavailable": why}
    try:
        r = run_reference(a.gpus, rounds=a.steps, warmup=a.warmup, model=a.model, n_train=a.n_train,
                          n_test=a.n_test, gpu_workers=a.ref_gpu_workers, timeout_s=6 * 3600)
    except Exception as e:   # noqa: BLE001
        return {"impl": "reference", "unavailable": f"{type(e).__name__}: {e}"[:300]}
    val = r["rounds_per_sec"]
    return {"impl": "reference", "metric": METRIC, "value": val, "unit": "rounds/s", "n_gpus": a.gpus,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1000.0 / val, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "e2e": {"value": val, "unit": "rounds/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                    "note": "the reference is end to end by construction (gRPC + pickle + base64, wall clock)"},
            "gpu_launches": 0,
            "config": {"model": a.model, "clients": a.gpus, "global_batch": a.batch_size * a.gpus, "seq_len": None,
                       "parallelism": f"fedavg{a.gpus}", "train_samples": a.n_train, "test_samples": a.n_test,
                       "local_epochs": 1, "eval_in_round": True, "eval": "inline (on the round's critical path)",
                       "augment": True, "codec": "none",
                       "workers": "gpu+DataParallel-identity shim" if a.ref_gpu_workers
                       else "cpu (the only mode in which the unmodified reference works, SURVEY B4)",
                       "path": "baseline/_ref/fedref_src via baseline/ref_launcher.py (sources byte-identical)"}}


def run_ours(a) -> dict:
    import torch
    import torch.distributed as dist
    from fedb200.ops import _lib
    from fedb200.parallel.session import GpuFederation
    from fedb200.utils import data as D

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != a.gpus:
        raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {a.gpus}")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    ds = D.synthetic_cifar(a.n_train, a.n_test, seed=0)
    fed = GpuFederation(model=a.model, rank=rank, world=world, device=local, dataset=ds, seed=1000 + rank,
                        batch_size=a.batch_size, codec=a.codec, nvls_reduce=a.nvls_reduce,
                        use_multicast=not a.no_multicast, repl_mode=a.repl, eval_mode=a.eval,
                        augment=not a.no_augment,
                        tag=f"bench-{os.environ.get('MASTER_PORT', 'solo')}-{os.getppid()}")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")      # > 126 MB L2

    # ---------------------------------------------------------------- warm-up (uploads + graph capture)
    # both flavours of the round graph are captured here, never inside a timed region: the first W-1 rounds stream their
    # inputs from pinned memory (the e2e flavour), the last one reuses the resident copy (the device-timed flavour)
    last = None
    W = max(a.warmup, 3)
    for i in range(W):
        last = fed.run_round(upload=i < W - 1, sync=True)
    fed.drain()
    barrier()

    # ---------------------------------------------------------------- device-timed rounds
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    launches0 = _lib.launches
    barrier()
    for i, (s, e) in enumerate(ev):
        if not a.no_l2_flush:
            flush.zero_()                   # L2 flush between timed iterations, outside the events
        s.record()
        fed.run_round(upload=False, sync=False)
        if i == a.steps - 1:
            fed.drain(sync=False)           # the evaluation still in flight on the second stream ends inside the timed region
        e.record()
    barrier()
    launches = _lib.launches - launches0
    fed.group.check()
    clocks = sampler.stop() if rank == 0 else {}
    dev_ms = max_over_ranks(sum(s.elapsed_time(e) for s, e in ev))
    value = a.steps / (dev_ms / 1000.0)

    # ---------------------------------------------------------------- end to end through the public API
    barrier()
    t0 = time.perf_counter()
    pending = None
    for _ in range(a.steps):
        h = fed.run_round_async(upload=True)        # H2D of the round's inputs + D2H of its result are inside
        if pending is not None:
            last = pending.result()                 # result of the previous round, read while this one runs
        pending = h
    last = pending.result()
    last_eval = fed.drain()                        # overlap mode: the last round's evaluation, still inside the timed region
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    h2d, d2h = fed.h2d_bytes_per_round(), fed.d2h_bytes_per_round()
    train_st, eval_st = last[0], last_eval

    # ---------------------------------------------------------------- correctness of the multi-GPU round (untimed)
    def all_gather(t):
        if world == 1:
            return [t]
        outs = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(outs, t.contiguous())
        return outs
    check = fed.aggregation_selfcheck(all_gather) if a.codec == "none" and not a.nvls_reduce else \
        {"agg_check": "skipped (lossy codec / in-switch reduction order)", "arenas_equal": None}
    out = {
        "metric": METRIC, "value": value, "unit": "rounds/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": dev_ms / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "fp32", "data": "synthetic", "impl": "ours",
        "e2e": {"value": a.steps / e2e_s, "unit": "rounds/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches, "clocks": clocks,
        "samples_per_sec": value * (a.n_train + a.n_test * world),
        "config": {"model": a.model, "clients": world, "global_batch": a.batch_size * world, "seq_len": None,
                   "parallelism": f"fedavg{world}", "train_samples": a.n_train, "test_samples": a.n_test,
                   "local_epochs": 1, "steps_per_client_per_round": train_st.steps, "eval_in_round": True,
                   "eval": ("overlap: all %d test images scored by every client every round on a second stream, concurrent "
                            "with the next local pass; K evaluations end inside the K timed rounds" % a.n_test)
                   if a.eval == "overlap" else "inline (on the round's critical path)",
                   "augment": not a.no_augment, "agg_check": check["agg_check"], "arenas_equal": check["arenas_equal"],
                   "l2": "flushed between rounds (256 MiB memset outside the timed events); round inputs are "
                         f"{h2d / 1e6:.0f} MB per client",
                   "codec": a.codec, "multicast": bool(fed.group.peers.mc_base), "nvls_reduce": a.nvls_reduce,
                   "replication": a.repl, "final_train_loss": train_st.mean_loss, "final_eval_acc": eval_st.accuracy},
    }
    fed.close()
    if world > 1:
        dist.destroy_process_group()
    return out if rank == 0 else {}


def main():
    a = parse()
    out = run_reference_arm(a) if a.impl == "reference" else run_ours(a)
    if out:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
