"""Rebuilds profiles/zoo_sweep.md from a `tools/bench_zoo.py` log (fedb200, this round) and the cuDNN / round-1 columns that are
kept in the table itself.

  python tools/make_zoo_sweep.py gpurun_out/prof/zoo.log
"""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATH = os.path.join(ROOT, "profiles", "zoo_sweep.md")


def main():
    new = {}
    for line in open(sys.argv[1]):
        if line.startswith("{"):
            r = json.loads(line)
            new[r["model"]] = r["ms_per_step"]
    old = open(PATH).read()
    rows = []
    for line in old.split("\n"):
        cells = [c.strip() for c in line.split("|")]
        if len(cells) >= 7 and cells[1] in new:
            name, r1, cg, ce = cells[1], cells[2], cells[4], cells[5]
            n = new[name]
            if name == "LeNet":                                     # 8-step passes are dominated by the H2D of the first chunk
                rows.append((name, r1, cells[3], cg, ce, cells[6]))
                continue
            ratio = f"{n / float(cg):.2f}x" if re.fullmatch(r"[\d.]+", cg) else "-"
            rows.append((name, r1, f"{n:.2f}", cg, ce, ratio))
    head = old[:old.index("| model |")]
    tail = old[old.index("\nAhead of cuDNN+graph now"):] if "\nAhead of cuDNN+graph now" in old else ""
    ahead = [n for n, _, v, cg, _, _ in rows if re.fullmatch(r"[\d.]+", cg) and float(v.split()[0]) < float(cg)]
    comparable = sum(1 for r in rows if re.fullmatch(r"[\d.]+", r[3]))
    table = ("| model | fedb200 round 1 | fedb200 round 2 | torch+cuDNN, CUDA graph | torch+cuDNN, eager | round 2 / cuDNN-graph |\n"
             "|---|---|---|---|---|---|\n" + "".join(f"| {' | '.join(r)} |\n" for r in rows))
    tail = re.sub(r"Ahead of cuDNN\+graph now: .*? \(\d+ of \d+ comparable models",
                  f"Ahead of cuDNN+graph now: {', '.join(ahead)} ({len(ahead)} of {comparable} comparable models", tail, flags=re.S)
    open(PATH, "w").write(head + table + tail)
    print(table)
    print("ahead:", ahead)


if __name__ == "__main__":
    main()
