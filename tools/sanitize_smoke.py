"""Small workload that touches every kernel family once, for `compute-sanitizer --tool memcheck|racecheck`:
  compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_smoke.py
LeNet step + eval, tcgen05 GEMM / implicit-GEMM conv (stride 1 and 2, grouped, im2col stem), depthwise, BN, pooling,
elementwise, loss, optimiser, codec, single-rank K-AGG.  No CUDA graphs (the sanitizer instruments plain launches)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedb200 import models  # noqa: E402
from fedb200.control.trainer import LocalTrainer, TrainerConfig  # noqa: E402
from fedb200.ops import cuda as K  # noqa: E402
from fedb200.ops import functional as F  # noqa: E402
from fedb200.utils import data as D  # noqa: E402

torch.manual_seed(0)
dev = "cuda:0"
ds = D.synthetic_cifar(128, 20, seed=0)
t = LocalTrainer(TrainerConfig(model="LeNet", device=dev, seed=0, use_cuda_graph=False), ds)
st = t.train(1, 0, 1)
ev = t.test()
print("lenet", st.steps, round(st.mean_loss, 4), ev.samples)


def conv(c, k, fs, stride, groups, hw, b=8):
    x = torch.randn(b, c, hw, hw, device=dev, requires_grad=True)
    w = (torch.randn(k, c // groups, fs, fs, device=dev) * 0.1).requires_grad_(True)
    y = F.conv2d(x, w, None, stride, fs // 2, groups)
    y.backward(torch.randn_like(y))
    return float(y.abs().sum())


print("conv", conv(64, 64, 3, 1, 1, 16), conv(64, 128, 3, 2, 1, 16), conv(64, 128, 1, 2, 1, 16), conv(48, 24, 1, 1, 1, 16),
      conv(128, 128, 3, 1, 4, 8), conv(3, 32, 3, 1, 1, 16), conv(32, 32, 3, 1, 32, 16), conv(24, 24, 5, 2, 24, 16), conv(20, 12, 3, 1, 1, 16))
m = models.build("MobileNet").to(dev).train()
x = torch.randn(8, 3, 32, 32, device=dev)
loss = F.cross_entropy(m(x), torch.randint(0, 10, (8,), device=dev))
loss.backward()
print("mobilenet", round(float(loss), 4))
g = models.build("GoogLeNet").to(dev).train()
loss = F.cross_entropy(g(x[:4]), torch.randint(0, 10, (4,), device=dev))
loss.backward()
print("googlenet", round(float(loss), 4))
n = 4096
cur, base, res = torch.randn(n, device=dev), torch.randn(n, device=dev), torch.zeros(n, device=dev)
q, sc = torch.zeros(n, dtype=torch.uint8, device=dev), torch.zeros(n // 32, device=dev)
K.codec_encode(cur, base, res, q, sc)
out = torch.empty(n, device=dev)
K.codec_decode(out, base, q, sc)
p, gr, mo = torch.randn(n, device=dev), torch.randn(n, device=dev), torch.zeros(n, device=dev)
K.sgd_step(p, gr, mo, 0.1, 0.9, 5e-4)
torch.cuda.synchronize()
print("SANITIZE_SMOKE_DONE")
