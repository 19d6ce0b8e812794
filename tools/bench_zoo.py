"""Local-training throughput of zoo models on one GPU through the public trainer (hand-written kernels, CUDA graph):
samples/s at batch 128 on synthetic CIFAR-shaped data.  Usage: python tools/bench_zoo.py MobileNet ResNet18 ..."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedb200.control.trainer import LocalTrainer, TrainerConfig  # noqa: E402
from fedb200.ops import _lib  # noqa: E402
from fedb200.utils import data as D  # noqa: E402

names = sys.argv[1:] or ["MobileNet", "ResNet18"]
steps = int(os.environ.get("ZOO_STEPS", "8"))
ds = D.synthetic_cifar(128 * steps, 256, seed=0)
for name in names:
    t = LocalTrainer(TrainerConfig(model=name, device="cuda:0", seed=0, use_cuda_graph=os.environ.get("ZOO_GRAPH", "1") == "1"), ds)
    t.train(1, 0, 1)                                   # warm-up (+ graph capture)
    torch.cuda.synchronize()
    l0 = _lib.launches
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    st = t.train(1, 0, 1)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b)
    print(json.dumps({"model": name, "steps": st.steps, "ms_per_step": ms / st.steps, "samples_per_s": st.samples / ms * 1e3,
                      "loss": st.mean_loss, "launches_per_step": (_lib.launches - l0) / max(st.steps, 1)}), flush=True)
    del t
    import gc
    gc.collect()
    torch.cuda.empty_cache()
