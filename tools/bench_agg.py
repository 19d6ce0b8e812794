"""FedAvg aggregation bandwidth sweep (BASELINE config 5): K-AGG two-shot (unicast P2P stores / NVLS multicast store /
NVLS in-switch reduce) against a plain NCCL all-reduce FedAvg, model sizes 1 MB .. 2 GB, N = WORLD_SIZE ranks.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29520 \
      tools/bench_agg.py [--max-mb 2048]

Every number is device-timed with CUDA events after warm-up, max over ranks.  "algbw" = model bytes / time.
Roofline: the busiest NVLink direction of a GPU is its INBOUND side - (N-1) shards arrive as answers to its reads and
(N-1) shards arrive as the peers' broadcast writes - so t >= 2 * (N-1)/N * model bytes / 770 GB/s (measured peer
copy, per direction); with the in-switch reduction the reads shrink to one shard: (1 + (N-1)) / N * bytes / 770.
The same bound applies to NCCL's reduce-scatter + all-gather.
Rank 0 prints a table and writes gpurun_out/agg_sweep_N<world>.json.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedb200.ops import cuda as K  # noqa: E402
from fedb200.parallel.arena import split_ranges  # noqa: E402
from fedb200.parallel.mem import PeerGroup, PeerLayout  # noqa: E402
from fedb200.parallel.session import dist_min_barrier  # noqa: E402

NVLINK_GBS = 770.0


def timed(fn, iters, world):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)          # ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=2048)
    ap.add_argument("--min-mb", type=float, default=1)
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rows = []
    mb = args.min_mb
    while mb <= args.max_mb:
        n = int(mb * (1 << 20) / 4) // 1024 * 1024
        layout = PeerLayout(n, 16, with_codec=False)
        pg = PeerGroup(rank, world, local, layout, f"agg-{os.environ.get('MASTER_PORT', 'solo')}-{int(mb * 100)}",
                       barrier=dist_min_barrier())
        me = pg.me
        me.client_f32().normal_()
        torch.cuda.synchronize()
        begin, end = split_ranges(n, world)[rank]
        bufs = pg.buffers
        common = dict(ctrl_local=me.ctrl_ptr, sync_mode=1, rank=rank, world=world, ctrl_peer=[b.ctrl_ptr for b in bufs])
        src = [b.client_f32_ptr() for b in bufs]
        aux0 = []          # the global model is a local copy on the primary (see FedGroup.aggregate)
        iters = 20 if mb <= 64 else (8 if mb <= 512 else 4)

        def p2p():
            K.fedavg_raw(src, aux0 + [b.client_f32_ptr() for b in bufs], begin, end, **common)

        def mcast():
            K.fedavg_raw(src, aux0, begin, end, mc_dst=pg.mc_client_f32_ptr(), **common)

        def nvls():
            K.fedavg_raw(src, aux0, begin, end, mc_dst=pg.mc_client_f32_ptr(), mc_src=pg.mc_client_f32_ptr(), **common)

        flat = torch.randn(n, device="cuda")

        def nccl():
            if world > 1:
                dist.all_reduce(flat)
            flat.div_(world)

        row = {"mb": mb, "n_f32": n, "world": world, "p2p_ms": timed(p2p, iters, world)}
        if pg.mc_base:
            row["mcast_ms"] = timed(mcast, iters, world)
            row["nvls_ms"] = timed(nvls, iters, world)
        row["nccl_ms"] = timed(nccl, iters, world)
        st = K.ctrl_field(me.ctrl(), "status")
        assert st == 0, f"flag wait failed: {st}"
        nbytes = n * 4
        if world > 1:
            roof_ms = 2 * nbytes * (world - 1) / world / (NVLINK_GBS * 1e9) * 1e3
            roof_nvls_ms = nbytes / (NVLINK_GBS * 1e9) * 1e3
        else:
            roof_ms = roof_nvls_ms = nbytes * 2 / 6.5459e12 * 1e3
        row["roofline_ms"], row["roofline_nvls_ms"] = roof_ms, roof_nvls_ms
        for k in ("p2p", "mcast", "nvls", "nccl"):
            if f"{k}_ms" in row:
                row[f"{k}_algbw_gbs"] = nbytes / row[f"{k}_ms"] / 1e6
                row[f"{k}_frac_roofline"] = (roof_nvls_ms if k == "nvls" else roof_ms) / row[f"{k}_ms"]
        best = min(row[f"{k}_ms"] for k in ("p2p", "mcast", "nvls") if f"{k}_ms" in row)
        row["best_vs_nccl"] = row["nccl_ms"] / best
        rows.append(row)
        if rank == 0:
            print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in row.items()}), flush=True)
        del flat
        if world > 1:
            dist.barrier()
        pg.close()
        torch.cuda.empty_cache()
        mb *= 4
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/agg_sweep_N{world}.json", "w") as f:
            json.dump(rows, f, indent=1)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
