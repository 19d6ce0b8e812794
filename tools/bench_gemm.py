"""tcgen05 GEMM throughput on the shapes the zoo produces (MobileNet / ResNet-50 pointwise convolutions at batch 128)
plus one large square problem; compares with torch.matmul (cuBLAS, TF32 allowed) as the library yardstick.
CUDA-event timing, 3 warm-ups, L2 flushed between timed iterations."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedb200.ops import cuda_gemm as G  # noqa: E402

torch.backends.cuda.matmul.allow_tf32 = True
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot / iters


rows = []
shapes = [("mobilenet pw 32->64 @32x32", 131072, 64, 32), ("mobilenet pw 128->128 @16x16", 32768, 128, 128),
          ("mobilenet pw 256->256 @8x8", 8192, 256, 256), ("mobilenet pw 512->512 @4x4", 2048, 512, 512),
          ("mobilenet pw 1024->1024 @2x2", 512, 1024, 1024), ("resnet50 pw 64->256 @32x32", 131072, 256, 64),
          ("resnet50 pw 256->64 @32x32", 131072, 64, 256), ("square 8192", 8192, 8192, 8192)]
only = sys.argv[1:] and [int(x) for x in sys.argv[1:]]
for idx, (name, M, N, K) in enumerate(shapes):
    if only and idx not in only:
        continue
    a, b = torch.randn(M, K, device="cuda"), torch.randn(N, K, device="cuda")
    d = torch.empty(M, N, device="cuda")
    ms = timeit(lambda: G.gemm_tf32(a, b, d, M, N, K, K, K, N))
    ms_lib = timeit(lambda: torch.matmul(a, b.t(), out=d))
    fl = 2.0 * M * N * K
    byt = 4.0 * (M * K + N * K + M * N)
    rows.append({"shape": name, "M": M, "N": N, "K": K, "ms": ms, "tflops": fl / ms / 1e9, "gbs": byt / ms / 1e6,
                 "cublas_ms": ms_lib, "cublas_tflops": fl / ms_lib / 1e9})
    print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in rows[-1].items()}), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/gemm_bench.json", "w"), indent=1)
