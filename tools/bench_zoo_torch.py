"""Library baseline for tools/bench_zoo.py: the REFERENCE's own model definitions (baseline/_ref/fedref_src/models, plain
torch.nn) trained with stock PyTorch on the same GPU - cuDNN / cuBLAS kernels, TF32 allowed, channels_last, the whole
step (forward, loss, backward, SGD) captured in one CUDA graph.  This is what "just move the reference to the GPU"
buys; none of fedb200's kernels run here.  Usage: python tools/bench_zoo_torch.py MobileNet ResNet18 ..."""
import importlib
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
names = sys.argv[1:] or ["MobileNet", "ResNet18"]
steps = int(os.environ.get("ZOO_STEPS", "8"))
torch.backends.cudnn.benchmark = True
torch.backends.cudnn.allow_tf32 = True
torch.backends.cuda.matmul.allow_tf32 = True
models = importlib.import_module("fedref_src.models")
dev = torch.device("cuda:0")
def run(name, channels_last):
    torch.manual_seed(0)
    fmt = torch.channels_last if channels_last else torch.contiguous_format
    if name.startswith("VGG"):
        ctor = lambda: models.VGG(name)                                            # noqa: E731
    elif name.startswith("ShuffleNetV2_"):
        ctor = lambda: models.ShuffleNetV2(net_size=float(name.split("_")[1]))     # noqa: E731
    else:
        ctor = getattr(models, name)
    net = ctor().to(dev).to(memory_format=fmt)
    opt = torch.optim.SGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4)
    crit = torch.nn.CrossEntropyLoss()
    x = torch.randn(128, 3, 32, 32, device=dev).contiguous(memory_format=fmt)
    y = torch.randint(0, 10, (128,), device=dev)

    def step():
        opt.zero_grad(set_to_none=False)
        loss = crit(net(x), y)
        loss.backward()
        opt.step()
        return loss

    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            step()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        step()
    b.record()
    torch.cuda.synchronize()
    eager = a.elapsed_time(b) / steps
    graphed = None
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            step()
        g.replay()
        torch.cuda.synchronize()
        a.record()
        for _ in range(steps):
            g.replay()
        b.record()
        torch.cuda.synchronize()
        graphed = a.elapsed_time(b) / steps
    except Exception as e:   # noqa: BLE001
        graphed = f"capture failed: {type(e).__name__}"
    print(json.dumps({"model": name, "impl": "torch+cudnn (reference model code)", "channels_last": channels_last,
                      "eager_ms_per_step": eager, "graph_ms_per_step": graphed}), flush=True)


for name in names:
    try:
        try:
            run(name, True)
        except RuntimeError:      # some reference models call .view() on the activations: NCHW only
            run(name, False)
    except Exception as e:        # noqa: BLE001 - e.g. the reference's ShuffleNetG constructors raise TypeError (SURVEY B-list)
        print(json.dumps({"model": name, "impl": "torch+cudnn (reference model code)", "error": f"{type(e).__name__}: {e}"[:160]}),
              flush=True)
    torch.cuda.empty_cache()
