"""Which kernel family takes every convolution of the zoo (forward / data gradient / weight gradient) at batch 128.
Pure shape logic: runs on the CPU.  Usage: python tools/conv_coverage.py [model ...]"""
import collections
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedb200 import models  # noqa: E402
from fedb200.ops import cuda_gemm as G  # noqa: E402
from fedb200.ops import cuda_nn as NN  # noqa: E402
from fedb200.ops.layers import Conv2d  # noqa: E402


def classify(shp):
    N, H, W, Cc, K, R, S, P, Q, sh, sw, ph, pw, groups = shp
    if NN._is_depthwise(shp):
        return ("depthwise",) * 3
    if groups > 1:
        gf, gd = G._group_geometry(shp, "fwd"), G._group_geometry(shp, "dgrad")
        Cg, Kg = Cc // groups, K // groups
        geo = G._group_geometry(shp, "wgrad") is not None
        return ("tc-grouped" if gf else "generic", "tc-grouped" if gd else "generic",
                "tc-grouped" if geo and ((Cg % 8 == 0 and Kg % 8 == 0) or groups <= G.MAX_SPLIT_GROUPS) else "generic")
    if G._small_reduction(shp):
        return ("tc-im2col", "-", "tc-im2col")
    f = d = w = "generic"
    st = G._strided(shp)
    if G._is_3x3(shp, Cc, K):
        f = "tc"
    elif st is not None and Cc % 4 == 0 and K % 4 == 0:
        f = "tc"
    elif G._is_pointwise(shp, 4, 1):
        f = "tc"
    if G._is_3x3(shp, K, Cc):
        d = "tc"
    elif st is not None and K % 4 == 0 and Cc % 4 == 0:
        d = "tc"
    elif G._is_pointwise(shp, 1, 4):
        d = "tc"
    if G._is_3x3(shp, Cc, K) or st is not None or G._is_pointwise(shp, 1, 1):
        w = "tc"
    return f, d, w


def conv_rows(name):
    """[(classification, forward FLOPs at batch 128, shape14)] for every convolution of a zoo model."""
    net = models.build(name).eval()
    rows = []

    def hook(mod, inp, out):
        x = inp[0]
        w = mod.weight
        shp = NN._conv_shape(torch.empty(128, *x.shape[1:], device="meta"), w, mod.stride, mod.padding, mod.groups)
        flops = 2 * 128 * shp[7] * shp[8] * shp[4] * (shp[3] // shp[13]) * shp[5] * shp[6]
        rows.append((classify(shp), flops, shp))

    hs = [m.register_forward_hook(hook) for m in net.modules() if isinstance(m, Conv2d)]
    with torch.no_grad():
        net(torch.zeros(2, 3, 32, 32))
    for h in hs:
        h.remove()
    return rows


def generic_share(name):
    """Share of the convolution FLOPs of a model that still runs on the generic fp32 kernels, per direction."""
    rows = conv_rows(name)
    tot = sum(r[1] for r in rows) or 1
    gen = collections.Counter()
    for (f, d, w), fl, shp in rows:
        for tag, v in (("fwd", f), ("dgrad", d), ("wgrad", w)):
            if v == "generic":
                gen[tag] += fl
    return {tag: gen[tag] / tot for tag in ("fwd", "dgrad", "wgrad")}


def main():
    names = sys.argv[1:] or sorted(models.REGISTRY)
    for name in names:
        rows = conv_rows(name)
        tot = sum(r[1] for r in rows) or 1
        gen = {k: v * tot for k, v in generic_share(name).items()}
        worst = [(fl, shp, c) for c, fl, shp in rows if "generic" in c]
        worst.sort(reverse=True)
        print(f"{name:18s} convs {len(rows):3d}  generic FLOP share fwd {100 * gen['fwd'] / tot:5.1f}%  dgrad {100 * gen['dgrad'] / tot:5.1f}%"
              f"  wgrad {100 * gen['wgrad'] / tot:5.1f}%" + (f"   e.g. C={worst[0][1][3]} K={worst[0][1][4]} R={worst[0][1][5]} "
              f"stride={worst[0][1][9]} groups={worst[0][1][13]} map={worst[0][1][1]} {worst[0][2]}" if worst else ""))


if __name__ == "__main__":
    main()
