import sys, torch
sys.path.insert(0, '/root/repo')
import torch.nn.functional as TF
from fedb200.ops import cuda_gemm as G
from fedb200.ops._lib import check, current_stream
for (B, C, K, HW, splits) in [(2, 64, 64, 8, 1), (2, 64, 64, 8, 2), (4, 64, 64, 32, 1), (4, 128, 64, 16, 4)]:
    torch.manual_seed(0)
    x = torch.randn(B, C, HW, HW).cuda().contiguous(memory_format=torch.channels_last)
    dy = torch.randn(B, K, HW, HW).cuda().contiguous(memory_format=torch.channels_last)
    w = torch.zeros(K, C, 3, 3, requires_grad=True)
    xr = x.detach().cpu().to(torch.bfloat16).float(); dyr = dy.detach().cpu().to(torch.bfloat16).float()
    y = TF.conv2d(xr, w, padding=1); y.backward(dyr)
    want = w.grad   # [K,C,3,3]
    dwk = torch.zeros(K * 9 * C, device="cuda")
    dy16, x16 = G.cast_bf16(dy), G.cast_bf16(x)
    check(G._f("fedb_conv3x3_wgrad_bf16")(dy16.data_ptr(), x16.data_ptr(), dwk.data_ptr(), B, HW, HW, C, K, splits, current_stream()))
    torch.cuda.synchronize()
    got = dwk.view(K, 9, C).permute(0, 2, 1).reshape(K, C, 3, 3).cpu()
    err = (got - want).abs().max() / want.abs().max()
    per_tap = [(float((got[:, :, t // 3, t % 3] - want[:, :, t // 3, t % 3]).abs().max() / want.abs().max())) for t in range(9)]
    ratio = float((got * want).sum() / (want * want).sum())
    print((B, C, K, HW, splits), "rel err", float(err), "ratio", ratio, "per tap", [round(v, 3) for v in per_tap])
