#!/usr/bin/env python
"""N-nccl baseline of BASELINE.md §3: "just put the reference on GPUs with the libraries".

Local training = the REFERENCE's own model code (baseline/_ref/fedref_src/models) on stock PyTorch: cuDNN/cuBLAS
kernels, TF32 allowed, one whole SGD step (forward, loss, backward, optimiser) captured in a CUDA graph and replayed
per batch.  Aggregation = ``torch.distributed.all_reduce`` (NCCL) over the flattened state, divided by N.  Evaluation =
the global model on the whole test set on every client.  Same round definition, shard rule, data shape and timing
method (CUDA events, max over ranks) as bench.py; none of fedb200's kernels run here.

  python baseline/nccl_fedavg.py --gpus 1 --steps 10 --warmup 3
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \\
         baseline/nccl_fedavg.py --gpus 8 --steps 10 --warmup 3
"""
import argparse
import importlib
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--model", default="LeNet")
    ap.add_argument("--n-train", type=int, default=50000)
    ap.add_argument("--n-test", type=int, default=10000)
    ap.add_argument("--batch-size", type=int, default=128)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == a.gpus
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    torch.backends.cudnn.benchmark = True
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True

    from fedb200.utils import data as D
    ds = D.synthetic_cifar(a.n_train, a.n_test, seed=0)
    bs = a.batch_size
    nb = (a.n_train + bs - 1) // bs
    mine = [b for b in range(nb) if (b + 1) % world == rank]                 # the reference's shard rule (N9)
    full = [b for b in mine if (b + 1) * bs <= a.n_train]                      # graph replays need a fixed batch size
    # inputs resident on the GPU, already normalised floats (generous to the baseline: no input pipeline at all)
    tx = D.normalize_batch(ds.train_x).to(dev).contiguous()
    ty = ds.train_y.to(dev)
    ex = D.normalize_batch(ds.test_x).to(dev).contiguous()
    ey = ds.test_y.to(dev)

    torch.manual_seed(1000 + rank)
    models = importlib.import_module("fedref_src.models")
    net = getattr(models, a.model)().to(dev)
    opt = torch.optim.SGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4)
    crit = torch.nn.CrossEntropyLoss()
    sx = torch.zeros(bs, 3, 32, 32, device=dev)
    sy = torch.zeros(bs, dtype=torch.long, device=dev)

    def step():
        opt.zero_grad(set_to_none=False)
        loss = crit(net(sx), sy)
        loss.backward()
        opt.step()

    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for _ in range(3):
            step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()

    state = [t for t in net.state_dict().values() if t.dtype.is_floating_point]
    flat = torch.zeros(sum(t.numel() for t in state), device=dev)

    def one_round():
        net.train()
        for b in full:
            sx.copy_(tx[b * bs:(b + 1) * bs])
            sy.copy_(ty[b * bs:(b + 1) * bs])
            graph.replay()
        if world > 1:
            torch.cat([t.reshape(-1) for t in state], out=flat)
            dist.all_reduce(flat)
            flat.div_(world)
            off = 0
            for t in state:
                t.copy_(flat[off:off + t.numel()].view_as(t))
                off += t.numel()
        net.eval()
        correct = torch.zeros((), device=dev)
        with torch.no_grad():
            for lo in range(0, a.n_test, 1000):
                correct += (net(ex[lo:lo + 1000]).argmax(1) == ey[lo:lo + 1000]).sum()
        return correct

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(a.warmup, 3)):
        one_round()
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    for s, e in ev:
        s.record()
        one_round()
        e.record()
    barrier()
    ms = sum(s.elapsed_time(e) for s, e in ev)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    if rank == 0:
        print(json.dumps({"impl": "N-nccl (stock PyTorch + cuDNN + CUDA-graph step + NCCL all-reduce)", "metric": "federated_rounds_per_sec",
                          "value": a.steps / (ms / 1e3), "unit": "rounds/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                          "ms_per_step": ms / a.steps, "dtype": "fp32 storage, TF32 math", "data": "synthetic, device resident",
                          "config": {"model": a.model, "clients": world, "steps_per_client_per_round": len(full),
                                     "train_samples": a.n_train, "test_samples": a.n_test}}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
