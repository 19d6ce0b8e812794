import sys, torch
sys.path.insert(0, '/root/repo')
import torch.nn.functional as TF
from fedb200.ops import functional as F
for (B, C, K, HW) in [(4, 32, 64, 32), (4, 64, 128, 16), (4, 128, 128, 16), (4, 128, 256, 8), (4, 256, 256, 8), (16, 64, 256, 32), (16, 256, 64, 32)]:
    torch.manual_seed(0)
    x = torch.randn(B, C, HW, HW); w = torch.randn(K, C, 1, 1) * 0.1
    xa, wa = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    xb, wb = x.clone().cuda().requires_grad_(True), w.clone().cuda().requires_grad_(True)
    ya = TF.conv2d(xa, wa); yb = F.conv2d(xb, wb)
    g = torch.randn_like(ya)
    ya.backward(g); yb.backward(g.cuda())
    rel = lambda a, b: float((a.cpu() - b).norm() / b.norm())
    print((B, C, K, HW), "fwd", rel(yb.detach(), ya.detach()), "dx", rel(xb.grad, xa.grad), "dw", rel(wb.grad, wa.grad))
