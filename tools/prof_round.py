"""Device-timed phases of federated rounds through the public API (GpuFederation(profile=True)): train / agg / repl /
repl_wait / eval milliseconds per rank, the NVLink bytes of the aggregation, the achieved GB/s and the fraction of the
primary -> backup replication that was hidden behind other work.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 \\
      tools/prof_round.py --model ResNet18 --rounds 4 --out profiles/round_profile_resnet18_N8.json
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedb200.parallel.session import GpuFederation  # noqa: E402
from fedb200.utils import data as D  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="ResNet18")
    ap.add_argument("--rounds", type=int, default=4)
    ap.add_argument("--n-train", type=int, default=50000)
    ap.add_argument("--n-test", type=int, default=10000)
    ap.add_argument("--eval", default="inline")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ds = D.synthetic_cifar(a.n_train, a.n_test, seed=0)
    fed = GpuFederation(model=a.model, rank=rank, world=world, device=local, dataset=ds, seed=1000 + rank, profile=True,
                        eval_mode=a.eval, tag=f"prof-{os.environ.get('MASTER_PORT', 'solo')}-{os.getppid()}")
    fed.run_round(upload=True, sync=True)    # warm-up: brings the inputs over ...
    for _ in range(2):                       # ... and captures the resident-input graphs that the measured rounds replay
        fed.run_round(upload=False, sync=True)
    fed.take_profile()
    recs = []
    for _ in range(a.rounds):
        fed.run_round(upload=False, sync=True)
        fed.group.before_local_pass()        # what the next round would wait for: the exposed part of the replication
        recs.append(fed.take_profile())
    mine = {k: sum(r.get(k, 0.0) for r in recs) / len(recs) for k in recs[-1] if k.endswith("_ms") or k.endswith("gbps")
            or k.endswith("fraction")}
    mine.update(rank=rank, nvlink_bytes=recs[-1]["nvlink_bytes"], arena_bytes=recs[-1]["arena_bytes"],
                is_primary=fed.group.rank == fed.group.primary_rank)
    allr = [None] * world
    if world > 1:
        dist.all_gather_object(allr, mine)
    else:
        allr = [mine]
    if rank == 0:
        out = {"model": a.model, "clients": world, "rounds_averaged": a.rounds, "eval_mode": a.eval,
               "multicast": bool(fed.group.peers.mc_base), "per_rank": allr,
               "max_over_ranks": {k: max(r.get(k, 0.0) for r in allr) for k in allr[0] if k.endswith("_ms")}}
        print(json.dumps(out))
        if a.out:
            with open(a.out, "w") as f:
                json.dump(out, f, indent=1)
    fed.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
