"""The communication kernels under ncu on ONE GPU (the profiling recipe forbids ncu on multi-rank commands).

Every "peer" arena of this run lives on the same GPU, so the loads and stores of K-AGG / K-REPL / K-PUSH / K-CODEC go to
local HBM instead of over NVLink: instruction mix, occupancy, loads in flight and the achieved memory throughput of the
very same SASS are visible, the link itself is not (the multi-GPU times and the bytes-over-NVLink arithmetic are in
tools/bench_agg.py / GpuFederation.take_profile).  One launch per kernel and size after a warm-up:

  python tools/prof_kagg.py [13 45 359]                                        # timing with CUDA events
  ncu --set full --clock-control none --import-source on -k regex:fedavg_kernel\\|replicate_kernel\\|push_copy_kernel\\|codec \\
      -o gpurun_out/kagg python tools/prof_kagg.py 45
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedb200.ops import cuda as K  # noqa: E402

DEV = "cuda:0"
N_PEERS = int(os.environ.get("KAGG_PEERS", "8"))


def timed(fn, iters=5):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    sizes = [float(x) for x in sys.argv[1:]] or [13.0, 45.0, 359.0]
    out = []
    for mb in sizes:
        n = int(mb * (1 << 20) / 4) // 1024 * 1024
        shard = n // N_PEERS // 4 * 4                     # what ONE rank of an N-rank two-shot K-AGG reduces and fans out
        src = [torch.randn(n, device=DEV) for _ in range(N_PEERS)]
        dst = [torch.empty(n, device=DEV) for _ in range(N_PEERS)]
        ctrl = K.new_ctrl(DEV)
        row = {"mb": mb, "n_f32": n, "peers": N_PEERS, "shard_f32": shard}

        def agg():     # fedavg_kernel<false,false,8>: pull the shard from every peer, ordered sum, push to every peer
            K.fedavg_raw([s.data_ptr() for s in src], [d.data_ptr() for d in dst], 0, shard, ctrl_local=ctrl.data_ptr())
        row["agg_ms"] = timed(agg)
        row["agg_gbs_local"] = 2 * N_PEERS * shard * 4 / row["agg_ms"] / 1e6

        q = [torch.zeros(n, dtype=torch.uint8, device=DEV) for _ in range(N_PEERS)]
        sc = [torch.zeros(n // 32, device=DEV) for _ in range(N_PEERS)]
        res = torch.zeros(n, device=DEV)
        base = torch.randn(n, device=DEV)

        def enc():     # codec_encode_kernel: fp8 block-scaled deltas + error feedback over the whole arena
            K.codec_encode(src[0], base, res, q[0], sc[0])
        row["encode_ms"] = timed(enc)
        for k in range(1, N_PEERS):
            K.codec_encode(src[k], base, res, q[k], sc[k])

        def agg_dec():  # fedavg_kernel<false,true,0>: fused decode of every peer's fp8 upload
            K.fedavg_raw([], [d.data_ptr() for d in dst], 0, shard, ctrl_local=ctrl.data_ptr(), src_q=[x.data_ptr() for x in q],
                         src_scale=[x.data_ptr() for x in sc], base=base.data_ptr())
        row["agg_decode_ms"] = timed(agg_dec)

        rep_ctrl = K.new_ctrl(DEV)

        def repl():    # K-REPL as the federation runs it: one source, global model + replica as destinations, version stamp
            K.fedavg_raw([src[0].data_ptr()], [dst[0].data_ptr(), dst[1].data_ptr()], 0, n, ctrl_local=ctrl.data_ptr(),
                         version=7, dst_ctrl=[0, rep_ctrl.data_ptr()])
        row["repl_ms"] = timed(repl)
        row["repl_gbs_local"] = 3 * n * 4 / row["repl_ms"] / 1e6

        def push():    # push_copy_kernel: one slice to N-1 destinations (unicast form)
            K.push_copy(src[0].data_ptr(), [d.data_ptr() for d in dst[1:]], shard * 4)
        row["push_ms"] = timed(push)
        out.append(row)
        print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in row.items()}), flush=True)
        del src, dst, q, sc, res, base
        torch.cuda.empty_cache()
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/prof_kagg.json", "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
