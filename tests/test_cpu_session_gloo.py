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
