#!/usr/bin/env python
"""Regenerates profiles/sass_evidence.md from the built library (runs on the CPU box: cuobjdump only).

Per kernel: instruction count and how often the mnemonics that prove the Blackwell path occur (the table of
B200_PROFILING.md: tcgen05.mma -> UTC*MMA, tcgen05.ld -> LDTM, TMA -> UTMALDG, mbarrier -> SYNCS, multimem -> *MC).
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "fedb200", "lib", "libfedb200.so")
WATCH = ("UTCHMMA", "UTCQMMA", "UTCOMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS",
         "LDGMC", "STGMC", "REDGMC", "HMMA", "FFMA", "LDS", "STS", "LDG", "STG", "SHFL", "BAR", "ATOMG", "REDG", "MEMBAR",
         "NANOSLEEP", "F2FP", "LDGSTS", "CCTL")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)", line)
        if m and cur:
            op = m.group(1)
            kernels[cur]["_total"] += 1
            base = op.split(".")[0]
            if base in WATCH:
                key = base
                if base in ("UTMALDG", "LDTM", "LDGMC", "STGMC", "SYNCS", "UTCBAR", "ATOMG", "REDG", "STG", "LDG") and "." in op:
                    key = ".".join(op.split(".")[:3]) if base in ("SYNCS", "LDGMC", "ATOMG") else ".".join(op.split(".")[:2])
                kernels[cur][key] += 1
    names = demangle(list(kernels))
    print("# SASS evidence (`cuobjdump -sass fedb200/lib/libfedb200.so`, sm_100a) - regenerate with `python tools/sass_evidence.py`\n")
    print("PTX names never appear in SASS: `tcgen05.mma` -> `UTCHMMA` (kind::tf32 / kind::f16), `tcgen05.ld` -> `LDTM`, "
          "`tcgen05.commit` -> `UTCBAR`, `tcgen05.alloc` -> `UTCATOMSWS`, TMA `cp.async.bulk.tensor` -> `UTMALDG.nD`, "
          "mbarrier -> `SYNCS`, `multimem.ld_reduce` -> `LDGMC...ADD`, `multimem.st` -> `STG` on the multicast VA "
          "(ptxas emits the plain store form; the address decides), `cp.async` -> `LDGSTS`.\n")
    print("| kernel | SASS instructions | watched mnemonics |")
    print("|---|---|---|")
    for k, c in kernels.items():
        tot = c.pop("_total", 0)
        nm = names.get(k, k).replace("(anonymous namespace)::", "").split("(")[0].replace("void ", "")
        hot = ", ".join(f"{m} x{n}" for m, n in sorted(c.items()))
        print(f"| `{nm}` | {tot} | {hot} |")
    # the multimem instructions, verbatim
    print("\nmultimem / tcgen05 instructions verbatim (first occurrences):\n\n```")
    seen = collections.Counter()
    for line in sass.splitlines():
        for tag in ("LDGMC", "UTCHMMA", "UTMALDG", "LDTM", "UTCBAR"):
            if tag in line and seen[tag] < 2:
                seen[tag] += 1
                print(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", "", line).strip())
    print("```")


if __name__ == "__main__":
    sys.exit(main())
