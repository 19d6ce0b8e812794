"""Small LeNet workload for ncu: a few eager training steps (4 kernels each) + one eval chunk."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedb200 import models
from fedb200.ops import cuda as K
from fedb200.parallel.arena import Arena
from fedb200.utils import data as D
torch.manual_seed(0)
m = models.LeNet().cuda()
arena = Arena.adopt(m, "cuda:0", with_optimizer_state=True)
plan = K.LeNetPlan(arena)
ds = D.synthetic_cifar(128 * 8, 2000, seed=0)
x, y = ds.train_x.cuda(), ds.train_y.cuda()
tx, ty = ds.test_x.cuda(), ds.test_y.cuda()
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 8
for s in range(steps):
    plan.train_step(x[(s % 8) * 128:(s % 8 + 1) * 128], y[(s % 8) * 128:(s % 8 + 1) * 128], step_seed=s)
plan.eval_batch(tx, ty)
torch.cuda.synchronize()
# timing of the eager loop with events (not under ncu)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for s in range(8):
        plan.train_step(x[s * 128:(s + 1) * 128], y[s * 128:(s + 1) * 128], step_seed=s)
g.replay(); torch.cuda.synchronize()
a.record()
for _ in range(20): g.replay()
b.record(); torch.cuda.synchronize()
print("graph: us per step", a.elapsed_time(b) * 1000 / (20 * 8))

if "--phases" in sys.argv:
    # per-phase SM cycles of CTA 0 of one sample-kernel launch (debug clock stamps, csrc/lenet.cu)
    import ctypes as C
    from fedb200.ops import _lib
    L = _lib.lib()
    L.fedb_lenet_phase_clocks.restype = C.c_int
    L.fedb_lenet_phase_clocks.argtypes = [C.c_void_p] * 3 + [C.c_int] + [C.c_void_p] * 3
    clk = torch.zeros(16, dtype=torch.int64, device="cuda")
    names = ["stage image+weights", "conv1+relu+pool", "conv2+relu+pool", "fc1/fc2/fc3 forward", "softmax-CE",
             "fc backward dh2, dh1", "fc backward dflat (+reduce)", "publish rows", "unpool2 + zero dp1", "conv2 wgrad",
             "conv2 dgrad scatter", "reduce dp1 + unpool1", "conv1 wgrad", "conv1 wgrad reduce + bias"]
    tot = None
    for rep in range(5):
        _lib.check(L.fedb_lenet_phase_clocks(C.byref(plan.desc), x[:128].data_ptr(), y[:128].data_ptr(), 128,
                                             plan.workspace.data_ptr(), clk.data_ptr(), None), "phase clocks")
        torch.cuda.synchronize()
        c = clk.cpu().tolist()
        d = [c[i + 1] - c[i] for i in range(14)]
        tot = d if tot is None else [min(a, b) for a, b in zip(tot, d)]
    s = sum(tot)
    print(f"phase cycles (min of 5), total {s} cycles = {s / 1.965e3:.1f} us at 1965 MHz")
    for nme, v in zip(names, tot):
        print(f"  {nme:28s} {v:7d}  {100 * v / s:5.1f} %")

if "--split" in sys.argv:
    # back-to-back sample kernels only (the debug entry launches just that kernel): step time - this = update + gaps
    import ctypes as C
    from fedb200.ops import _lib
    L = _lib.lib()
    L.fedb_lenet_phase_clocks.restype = C.c_int
    L.fedb_lenet_phase_clocks.argtypes = [C.c_void_p] * 3 + [C.c_int] + [C.c_void_p] * 3
    clk = torch.zeros(16, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream

    def only_sample():
        for s in range(8):
            _lib.check(L.fedb_lenet_phase_clocks(C.byref(plan.desc), x[s * 128:(s + 1) * 128].data_ptr(),
                                                 y[s * 128:(s + 1) * 128].data_ptr(), 128, plan.workspace.data_ptr(),
                                                 clk.data_ptr(), torch.cuda.current_stream().cuda_stream), "phase clocks")
    only_sample(); torch.cuda.synchronize()
    g2 = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g2):
        only_sample()
    g2.replay(); torch.cuda.synchronize()
    a.record()
    for _ in range(20): g2.replay()
    b.record(); torch.cuda.synchronize()
    print("graph: us per sample kernel alone", a.elapsed_time(b) * 1000 / (20 * 8))
