import sys, torch
sys.path.insert(0, '/root/repo')
from fedb200.ops import cuda as K
n = 4096
dev = "cuda:0"
base = torch.randn(n, device=dev)
curs = [base + 0.01*torch.randn(n, device=dev) for _ in range(2)]
qs, ss = [], []
for c in curs:
    q = torch.zeros(n, dtype=torch.uint8, device=dev); s = torch.zeros(n//32, device=dev); r = torch.zeros(n, device=dev)
    K.codec_encode(c, base, r, q, s); qs.append(q); ss.append(s)
i64 = [torch.full((16,), 190, dtype=torch.int64, device=dev), torch.full((16,), 193, dtype=torch.int64, device=dev)]
out = torch.zeros(n, device=dev); oi = torch.zeros(16, dtype=torch.int64, device=dev)
ctrl = K.new_ctrl(dev)
K.fedavg_raw([c.data_ptr() for c in curs], [out.data_ptr()], 0, n, ctrl_local=ctrl.data_ptr(), src_q=[q.data_ptr() for q in qs],
             src_scale=[s.data_ptr() for s in ss], base=base.data_ptr(), src_i64=[t.data_ptr() for t in i64], dst_i64=[oi.data_ptr()], n_i64=16)
torch.cuda.synchronize()
print("i64", oi.tolist()[:4], "f32 err", float((out - (curs[0]+curs[1])/2).abs().max()))
