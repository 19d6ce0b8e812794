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
