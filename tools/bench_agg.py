"""FedAvg aggregation bandwidth sweep (BASELINE config 5): K-AGG two-shot (unicast P2P stores / NVLS multicast store /
NVLS in-switch reduce) against a plain NCCL all-reduce FedAvg, model sizes 1 MB .. 2 GB, N = WORLD_SIZE ranks.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29520 \
      tools/bench_agg.py [--max-mb 2048]

Every number is device-timed with CUDA events after warm-up, max over ranks.  "algbw" = model bytes / time.
Roofline: the busiest NVLink direction of a GPU is its INBOUND side - (N-1) shards arrive as answers to its reads and
(N-1) shards arrive as the peers' broadcast writes - so t >= 2 * (N-1)/N * model bytes / B; with the in-switch reduction
the reads shrink to one shard: (1 + (N-1)) / N * bytes / B.  The same bound applies to NCCL's reduce-scatter + all-gather.
ONE denominator is used for the headline fraction: B = 900 GB/s, the NVLink 5 per-direction figure BASELINE.md states
(`frac_roofline`).  Next to it `frac_of_measured_copy` uses the peer-copy bandwidth measured in this very run (every rank
pushes 256 MB to its ring neighbour at the same time with the K-PUSH kernel; recorded as `measured_copy_gbs`).
The NCCL arm is timed twice: all_reduce(SUM) followed by a division pass, and all_reduce(AVG) in one call; `best_vs_nccl`
compares against the FASTER of the two.
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

NVLINK_GBS = 900.0      # NVLink 5, per direction


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
    ap.add_argument("--sizes-mb", type=float, nargs="*", default=None,
                    help="explicit sizes; default: powers of four plus the zoo's arenas (13 = MobileNet, 45 = ResNet-18, 359 = ResNeXt29)")
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rows = []
    # ---- peer-copy bandwidth of this box, this run: every rank pushes 256 MB to its ring neighbour concurrently
    measured_copy_gbs = None
    if world > 1:
        nb = 256 << 20
        pgc = PeerGroup(rank, world, local, PeerLayout(nb // 4, 16, with_codec=False),
                        f"aggc-{os.environ.get('MASTER_PORT', 'solo')}", use_multicast=False, barrier=dist_min_barrier())
        pgc.me.client_f32().normal_()
        dst = pgc.buffers[(rank + 1) % world].client_f32_ptr()
        srcp = torch.randn(nb // 4, device="cuda")
        ms = timed(lambda: K.push_copy(srcp.data_ptr(), [dst], nb), 10, world)
        measured_copy_gbs = nb / ms / 1e6
        dist.barrier()
        pgc.close()
        del srcp
        torch.cuda.empty_cache()
    sizes = args.sizes_mb
    if not sizes:
        sizes, mb = [], args.min_mb
        while mb <= args.max_mb:
            sizes.append(mb)
            mb *= 4
        sizes = sorted(set(sizes + [s for s in (13.0, 45.0, 359.0) if args.min_mb <= s <= args.max_mb]))
    for mb in sizes:
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

        def nccl_avg():
            if world > 1:
                dist.all_reduce(flat, op=dist.ReduceOp.AVG)

        row = {"mb": mb, "n_f32": n, "world": world, "p2p_ms": timed(p2p, iters, world)}
        if pg.mc_base:
            row["mcast_ms"] = timed(mcast, iters, world)
            row["nvls_ms"] = timed(nvls, iters, world)
        row["nccl_ms"] = timed(nccl, iters, world)
        row["nccl_avg_ms"] = timed(nccl_avg, iters, world)
        row["measured_copy_gbs"] = measured_copy_gbs
        st = K.ctrl_field(me.ctrl(), "status")
        assert st == 0, f"flag wait failed: {st}"
        nbytes = n * 4
        if world > 1:
            roof_ms = 2 * nbytes * (world - 1) / world / (NVLINK_GBS * 1e9) * 1e3
            roof_nvls_ms = nbytes / (NVLINK_GBS * 1e9) * 1e3
        else:
            roof_ms = roof_nvls_ms = nbytes * 2 / 6.5459e12 * 1e3
        row["roofline_ms"], row["roofline_nvls_ms"] = roof_ms, roof_nvls_ms
        row["roofline_gbs"] = NVLINK_GBS
        for k in ("p2p", "mcast", "nvls", "nccl", "nccl_avg"):
            if f"{k}_ms" in row:
                row[f"{k}_algbw_gbs"] = nbytes / row[f"{k}_ms"] / 1e6
                row[f"{k}_frac_roofline"] = (roof_nvls_ms if k == "nvls" else roof_ms) / row[f"{k}_ms"]
                if measured_copy_gbs:
                    row[f"{k}_frac_of_measured_copy"] = row[f"{k}_frac_roofline"] * NVLINK_GBS / measured_copy_gbs
        best = min(row[f"{k}_ms"] for k in ("p2p", "mcast", "nvls") if f"{k}_ms" in row)
        best_exact = min(row[f"{k}_ms"] for k in ("p2p", "mcast") if f"{k}_ms" in row)     # bit-reproducible paths only
        nccl_best = min(row["nccl_ms"], row["nccl_avg_ms"])
        row["best_vs_nccl"] = nccl_best / best
        row["deterministic_vs_nccl"] = nccl_best / best_exact
        rows.append(row)
        if rank == 0:
            print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in row.items()}), flush=True)
        del flat
        if world > 1:
            dist.barrier()
        pg.close()
        torch.cuda.empty_cache()
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/agg_sweep_N{world}.json", "w") as f:
            json.dump(rows, f, indent=1)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
