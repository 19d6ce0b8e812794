#!/usr/bin/env python
"""Turns .ncu-rep captures (gpurun_out/, scratch) into the tracked summaries under profiles/.

  python tools/ncu_summary.py gpurun_out/lenet_sample.ncu-rep [...]  > profiles/ncu_<name>.md

Per report: one table row per captured launch with the metrics the B200 profiling recipe names, plus the
ten hottest source lines (by warp-level executed instructions and by stall samples) when the report holds
source info (-lineinfo + --import-source on).  Runs on the CPU box: only `ncu -i` is needed.
"""
from __future__ import annotations

import csv
import io
import os
import subprocess
import sys

METRICS = [
    ("time", "gpu__time_duration.sum"),
    ("grid", "launch__grid_size"),
    ("block", "launch__block_size"),
    ("regs", "launch__registers_per_thread"),
    ("dyn_smem", "launch__shared_mem_per_block_dynamic"),
    ("dram_rd", "dram__bytes_read.sum"),
    ("dram_wr", "dram__bytes_write.sum"),
    ("dram_%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("sm_%", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("issue_%", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    ("fma_%", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
    ("lsu_%", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
    ("tensor_%", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
    ("warps_active_%", "sm__warps_active.avg.pct_of_peak_sustained_active"),
]


FILTER = os.environ.get("NCU_KERNEL", "")       # e.g. NCU_KERNEL=regex:update_kernel restricts every page to it


def ncu(args):
    extra = ["--kernel-name", FILTER, "--launch-count", "1"] if FILTER else []
    return subprocess.run(["ncu"] + args + extra, capture_output=True, text=True).stdout


def raw_rows(rep):
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    rows = [r for r in rows if len(r) > 10]
    if not rows:
        return [], [], []
    return rows[0], rows[1], rows[2:]


def short(name: str) -> str:
    name = name.replace("void ", "").replace("<unnamed>::", "")
    return name.split("(")[0]


def fmt(v: str) -> str:
    try:
        f = float(v.replace(",", ""))
    except ValueError:
        return v
    return f"{f:.0f}" if abs(f) >= 100 else f"{f:.2f}".rstrip("0").rstrip(".")


def source_hot(rep, top=12):
    """CUDA source lines with their aggregated SASS metrics (the `cuda,sass` view: per file a header row, then one
    row per source line followed by its SASS rows, which have an empty line number)."""
    out = ncu(["-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"])
    items, hdr, fname = [], None, ""
    for r in csv.reader(io.StringIO(out)):
        if len(r) == 2 and r[0] == "File Path":
            fname = os.path.basename(r[1])
        elif r and r[0] == "Line No":
            hdr = r
        elif hdr and len(r) == len(hdr) and r[0].isdigit():
            try:
                inst = float(r[hdr.index("Instructions Executed")].replace(",", "") or 0)
                samp = float(r[hdr.index("# Samples")].replace(",", "") or 0)
            except ValueError:
                continue
            if inst > 0 or samp > 0:
                items.append((inst, samp, f"{fname}:{r[0]}  {r[1].strip()}"))
    tot_i = sum(i for i, _, _ in items) or 1.0
    tot_s = sum(s for _, s, _ in items) or 1.0
    items.sort(key=lambda t: -t[1])
    return [(100 * i / tot_i, 100 * s / tot_s, src[:120]) for i, s, src in items[:top]]


def main():
    for rep in sys.argv[1:]:
        hdr, units, rows = raw_rows(rep)
        print(f"### `{os.path.basename(rep)}`\n")
        if not rows:
            print("(no launches in the report)\n")
            continue
        k = hdr.index("Kernel Name")
        cols = [(lbl, hdr.index(m)) for lbl, m in METRICS if m in hdr]
        print("| kernel | " + " | ".join(lbl for lbl, _ in cols) + " |")
        print("|---|" + "---|" * len(cols))
        for r in rows:
            print(f"| `{short(r[k])}` | " + " | ".join(fmt(r[i]) for _, i in cols) + " |")
        u = {lbl: units[i] for lbl, i in cols}
        print(f"\nunits as reported by ncu: time {u.get('time', '')}, dram read/write {u.get('dram_rd', '')}/"
              f"{u.get('dram_wr', '')}, smem {u.get('dyn_smem', '')}\n")
        hot = source_hot(rep)
        if hot:
            print("hottest source lines (share of executed warp instructions / of stall samples):\n")
            print("| inst % | samples % | source |")
            print("|---|---|---|")
            for i, s, src in hot:
                print(f"| {i:.1f} | {s:.1f} | `{src.replace('|', '/')}` |")
            print()


if __name__ == "__main__":
    main()
