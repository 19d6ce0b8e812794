"""Multi-GPU correctness of the peer plane; launch with torchrun (one rank per GPU):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      tools/multigpu_check.py

Checks, on every rank: peer mappings work, one K-AGG launch per rank gives the bit-exact FedAvg of all
ranks' arenas in every client arena, in the primary's global model and in the backup's replica (fused
and overlapped K-REPL), the int64 counters follow N3, multicast (NVLS) store path when available,
in-switch reduction (allclose), the fp8 delta codec (bounded error, error feedback), a shrunken live set.
"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedb200.parallel.fedavg import fedavg_flat_cpu  # noqa: E402
from fedb200.parallel.session import GpuFederation  # noqa: E402
from fedb200.utils import data as D  # noqa: E402


def gather(t):
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t.contiguous())
    return [o.cpu() for o in out]


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ds = D.synthetic_cifar(128 * 2 * world + 40, 500, seed=0)
    report = {}
    for mode in os.environ.get("MG_MODES", "p2p,fused-repl,nvls-reduce,fp8").split(","):
        kw = dict(repl_mode="fused" if mode == "fused-repl" else "overlap", nvls_reduce=mode == "nvls-reduce",
                  codec="fp8" if mode == "fp8" else "none")
        fed = GpuFederation(model=os.environ.get("MG_MODEL", "MobileNet" if mode != "fp8" else "LeNet"), rank=rank, world=world, device=local,
                            dataset=ds, seed=100 + rank, tag=f"mg-{mode}-{os.environ['MASTER_PORT']}", evaluate=False, **kw)
        man = fed.trainer.arena.manifest
        # make the counters differ per rank (N3) and the weights differ (different seeds already)
        fed.trainer.arena.i64[:man.n_i64] = 190 + 3 * rank
        if mode == "fp8":
            # round 0 is lossless by design (it establishes the common base); then every rank "trains"
            fed.group.aggregate()
            torch.cuda.synchronize()
            dist.barrier()
            fed.trainer.arena.f32[:man.n_f32] += 0.01 * torch.randn(man.n_f32, device="cuda")
            fed.trainer.arena.i64[:man.n_i64] = 190 + 3 * rank
        torch.cuda.synchronize()
        dist.barrier()
        pre_f = gather(fed.trainer.arena.f32[:man.n_f32])
        pre_i = gather(fed.trainer.arena.i64[:man.n_i64])
        fed.group.aggregate()
        torch.cuda.synchronize()
        fed.group.check()
        dist.barrier()
        torch.cuda.synchronize()
        want_f, want_i = torch.empty(man.n_f32), torch.empty(man.n_i64, dtype=torch.int64)
        fedavg_flat_cpu(pre_f, pre_i, None, want_f, want_i)
        got_f, got_i = fed.trainer.arena.f32[:man.n_f32].cpu(), fed.trainer.arena.i64[:man.n_i64].cpu()
        glob = fed.group.global_arena.f32[:man.n_f32].cpu()
        if mode in ("p2p", "fused-repl"):
            assert torch.equal(got_f, want_f), f"{mode}: client arena differs from the oracle"
            assert torch.equal(glob, want_f), f"{mode}: global model differs"
        elif mode == "nvls-reduce":
            if fed.group.peers.mc_base:
                torch.testing.assert_close(got_f, want_f, rtol=1e-5, atol=1e-6)
        else:   # fp8 deltas: bounded error against the exact mean (e4m3: <= 2^-4 of a block's amax ~ 0.04*2^-4)
            err = float((got_f - want_f).abs().max())
            report["fp8_max_abs_err"] = err
            assert err < 5e-3, f"fp8 error {err}"
            assert torch.equal(glob, got_f), "fp8: global model differs from the clients'"
        assert torch.equal(got_i, want_i), f"{mode}: int64 counters {got_i.tolist()} vs {want_i.tolist()} report={report}"
        if world > 1 and mode != "fp8":
            rep = fed.group.replica_arena.f32[:man.n_f32].cpu()
            torch.cuda.synchronize()
            fed.group.side_stream.synchronize()
            dist.barrier()
            rep = fed.group.replica_arena.f32[:man.n_f32].cpu()
            if mode != "nvls-reduce":
                assert torch.equal(rep, want_f), f"{mode}: backup replica differs"
        report[mode] = {"multicast": bool(fed.group.peers.mc_base), "mc_error": fed.group.peers.multicast_error}
        # all ranks agree bit for bit after the round
        allf = gather(fed.trainer.arena.f32[:man.n_f32])
        assert all(torch.equal(allf[0], a) for a in allf), f"{mode}: ranks disagree"
        # a full round through the public API (train + agg + eval) twice
        fed.evaluate = True
        tr, ev = fed.run_round()
        tr, ev = fed.run_round()
        assert tr.steps >= 1 and ev.samples == 500
        if world > 2 and mode == "p2p":
            live = list(range(world - 1))
            fed.group.set_live(live)
            pre_f = gather(fed.trainer.arena.f32[:man.n_f32])
            pre_i = gather(fed.trainer.arena.i64[:man.n_i64])
            fed.group.aggregate()
            torch.cuda.synchronize()
            dist.barrier()
            if rank in live:
                fedavg_flat_cpu([pre_f[r] for r in live], [pre_i[r] for r in live], None, want_f, want_i)
                assert torch.equal(fed.trainer.arena.f32[:man.n_f32].cpu(), want_f), "live-only mean"
            fed.group.set_live(list(range(world)))
        dist.barrier()
        fed.close()
        dist.barrier()
    # ---- evaluation-set all-gather over NVLink (K-PUSH): every rank uploads 1/N of the test images and pushes its slice
    # to all peers; the evaluation must give, on every rank, exactly what a full local upload gives
    for mc in (True, False):
        fed = GpuFederation(model="LeNet", rank=rank, world=world, device=local, dataset=ds, seed=100 + rank,
                            tag=f"mg-tex{int(mc)}-{os.environ['MASTER_PORT']}", evaluate=True, use_multicast=mc)
        assert fed.exchange is not None and fed.exchange.enabled()
        for r in range(4):                                   # both halves of the double buffer, live and as graph replays
            tr, ev = fed.run_round(upload=True)
            assert ev.samples == 500
        torch.cuda.synchronize()
        dist.barrier()
        h2d = fed.h2d_bytes_per_round()
        ev_local = fed.engine.eval_pass(upload=True, sync=True)      # full local upload into the current half
        assert (ev.correct, ev.loss_sum) == (ev_local.correct, ev_local.loss_sum), (ev, ev_local)
        allc = gather(torch.tensor([float(ev.correct), ev.loss_sum], device="cuda"))
        assert all(torch.equal(allc[0], c) for c in allc), "ranks disagree on the evaluation of the same global model"
        shard = fed.trainer.shard(rank, world)
        n_shard = sum(min(128, ds.train_x.shape[0] - b * 128) for b in shard)
        lo, hi = fed.exchange.rows(rank)
        assert h2d == n_shard * (3072 + 8) + (hi - lo) * 3072 + 500 * 8, (h2d, n_shard, lo, hi)
        report[f"exchange-mc{int(mc)}"] = {"multicast": bool(fed.exchange.peers.mc_base)}
        dist.barrier()
        fed.close()
        dist.barrier()
    if rank == 0:
        print("MULTIGPU_CHECK_OK", world, report)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
