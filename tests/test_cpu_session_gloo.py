"""CpuFederation under torch.distributed/gloo with world_size > 1 (runs on a box without GPUs): the rank-symmetric round of
fedb200/parallel/cpu_session.py must equal a sequential emulation of the reference's round loop (src/server.py:113-179)
with the oracle FedAvg - bit for bit, on every rank - including a round in which one rank is not live."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fedb200.control.trainer import LocalTrainer, TrainerConfig
from fedb200.parallel.fedavg import fedavg_state_dicts
from fedb200.utils import data as D

WORLD = 3
PLAN = [[0, 1, 2], [0, 2], [0, 1, 2]]          # live set of every round


def _dataset():
    return D.synthetic_cifar(128 * 7 + 30, 90, seed=11)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, port, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    from fedb200.parallel.cpu_session import CpuFederation
    fed = CpuFederation(model="LeNet", dataset=_dataset(), seed=40 + rank)
    log = []
    for live in PLAN:
        fed.set_live(live)
        tr, ev = fed.run_round()
        log.append((tr.steps, tr.samples, ev.samples, ev.correct, ev.loss_sum))
    torch.save({"state": fed.state_dict(), "log": log}, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_rounds_equal_the_sequential_oracle(tmp_path):
    mp.start_processes(_worker, args=(_free_port(), str(tmp_path)), nprocs=WORLD, join=True, start_method="spawn")
    got = [torch.load(tmp_path / f"rank{r}.pt", weights_only=False) for r in range(WORLD)]

    threads = torch.get_num_threads()
    torch.set_num_threads(1)                    # same summation order as the single-threaded workers
    try:
        _compare(got)
    finally:
        torch.set_num_threads(threads)


def _compare(got):
    ds = _dataset()
    trainers = [LocalTrainer(TrainerConfig(model="LeNet", device="cpu", seed=40 + r), ds) for r in range(WORLD)]
    want_log = [[] for _ in range(WORLD)]
    for live in PLAN:
        stats = {r: trainers[r].train(1, live.index(r), WORLD) for r in live}       # dense rank, world = total (N9)
        glob = fedavg_state_dicts([trainers[r].state_dict() for r in live])
        for r in range(WORLD):
            if r in live:
                trainers[r].load_state_dict(glob)
                ev = trainers[r].test()
                want_log[r].append((stats[r].steps, stats[r].samples, ev.samples, ev.correct, ev.loss_sum))
            else:
                want_log[r].append((0, 0, 0, 0, 0.0))
    for r in range(WORLD):
        assert [tuple(x) for x in got[r]["log"]] == want_log[r], r
        want = trainers[r].state_dict()
        for k, v in got[r]["state"].items():
            assert torch.equal(v, want[k]), (r, k)
    # after the last (all-live) round every rank holds the same global model
    for k, v in got[0]["state"].items():
        assert torch.equal(v, got[1]["state"][k]) and torch.equal(v, got[2]["state"][k])


def _worker_fp8(rank, port, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    from fedb200.parallel.cpu_session import CpuFederation
    fed = CpuFederation(model="LeNet", dataset=_dataset(), seed=40 + rank, codec="fp8", lr=0.05)
    sent = []
    for _ in range(3):
        fed.run_round()
        sent.append(fed.bytes_sent_last)
    torch.save({"state": fed.state_dict(), "sent": sent, "residual": fed._residual}, os.path.join(out_dir, f"fp8-{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_rounds_with_the_fp8_codec(tmp_path):
    """codec="fp8" on the CPU plane: round 1 goes out lossless (different random inits, N6), the following rounds ship one
    byte + 1/32 scale per weight; every rank ends with the SAME model, close to the lossless run of the same seeds, and keeps
    the dropped part as its own residual."""
    mp.start_processes(_worker_fp8, args=(_free_port(), str(tmp_path)), nprocs=WORLD, join=True, start_method="spawn")
    got = [torch.load(tmp_path / f"fp8-{r}.pt", weights_only=False) for r in range(WORLD)]
    n_f32 = sum(v.numel() for v in got[0]["state"].values())
    assert got[0]["sent"][0] >= 4 * n_f32                                   # lossless first round
    assert got[0]["sent"][1] == got[0]["sent"][2] < 0.3 * got[0]["sent"][0]  # then ~ 9 bits per weight
    for k, v in got[0]["state"].items():
        assert torch.equal(v, got[1]["state"][k]) and torch.equal(v, got[2]["state"][k]), k
    assert all(float(g["residual"].abs().max()) > 0 for g in got)
    # lossless run of the same seeds, sequentially
    ds = _dataset()
    threads = torch.get_num_threads()
    torch.set_num_threads(1)
    try:
        trainers = [LocalTrainer(TrainerConfig(model="LeNet", device="cpu", seed=40 + r, lr=0.05), ds) for r in range(WORLD)]
        for _ in range(3):
            for r in range(WORLD):
                trainers[r].train(1, r, WORLD)
            glob = fedavg_state_dicts([t.state_dict() for t in trainers])
            for t in trainers:
                t.load_state_dict(glob)
    finally:
        torch.set_num_threads(threads)
    want = trainers[0].state_dict()
    num = sum(float((got[0]["state"][k].float() - want[k].float()).pow(2).sum()) for k in want)
    den = sum(float(want[k].float().pow(2).sum()) for k in want)
    assert (num / den) ** 0.5 < 2e-2, (num / den) ** 0.5
