import sys, torch
sys.path.insert(0, '/root/repo')
from fedb200 import models
import torch.nn.functional as TF
from fedb200.ops import functional as F
name = sys.argv[1]
torch.manual_seed(6)
ref = models.build(name); gpu = models.build(name); gpu.load_state_dict(ref.state_dict()); gpu = gpu.to('cuda:0')
ref.train(); gpu.train()
x, y = torch.randn(4,3,32,32), torch.randint(0,10,(4,))
acts_r, acts_g = {}, {}
def hook(store):
    def mk(n):
        def h(m, i, o):
            if isinstance(o, torch.Tensor): store[n] = o.detach().float().cpu()
        return h
    return mk
for n, m in ref.named_modules(): m.register_forward_hook(hook(acts_r)(n))
for n, m in gpu.named_modules(): m.register_forward_hook(hook(acts_g)(n))
la = TF.cross_entropy(ref(x), y); lb = F.cross_entropy(gpu(x.cuda()), y.cuda())
la.backward(); lb.backward()
print("loss", float(la), float(lb))
cnt = 0
for n in acts_r:
    if n in acts_g:
        a, b = acts_r[n], acts_g[n]
        err = float((a-b).abs().max()) / (float(a.abs().max())+1e-9)
        if err > 1e-4 and cnt < 8:
            print("FWD", n, tuple(a.shape), err); cnt += 1
gp = dict(gpu.named_parameters())
rows = []
for k, p in ref.named_parameters():
    if p.grad is None: continue
    g = gp[k].grad.detach().cpu()
    rows.append((float((g-p.grad).abs().max())/(float(p.grad.abs().max())+1e-6), k, tuple(p.shape)))
rows.sort(reverse=True)
for r in rows[:8]: print("GRAD", r)
