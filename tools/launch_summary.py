"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name (shares, not absolutes).

  python tools/launch_summary.py [--last-step N_STEPS] file.csv ...

--last-step N: the run consisted of N equal steps; only the launches of the last one are counted (steady state: the first
step of an eager run also packs filters one by one and sizes workspaces)."""
import collections
import csv
import sys

args = sys.argv[1:]
steps = 0
if args and args[0] == "--last-step":
    steps = int(args[1])
    args = args[2:]
for path in args:
    rows = list(csv.reader(open(path)))
    hdr = [r for r in rows if "Kernel Name" in r][0]
    k, v = hdr.index("Kernel Name"), hdr.index("Metric Value")
    launches = []
    for r in rows:
        if len(r) == len(hdr) and r[v].replace(".", "").isdigit():
            launches.append((r[k].split("(")[0].replace("void ", "").replace("<unnamed>::", "")[-70:], float(r[v])))
    if steps > 1:
        # the last step starts at the last launch of the step's first kernel (the input pipeline's prepare_kernel)
        starts = [i for i, (n, _) in enumerate(launches) if "prepare_kernel" in n]
        launches = launches[starts[-1] - 1 if starts and starts[-1] > 0 else 0:]
    agg, cnt = collections.Counter(), collections.Counter()
    for n, t in launches:
        agg[n] += t
        cnt[n] += 1
    tot = sum(agg.values())
    try:
        print(f"## {path}: {sum(cnt.values())} launches, {tot / 1e3:.1f} us in kernels" + (" (last step)" if steps > 1 else ""))
        print("| kernel | launches | us | share |\n|---|---|---|---|")
        for n, t in agg.most_common(24):
            print(f"| `{n}` | {cnt[n]} | {t / 1e3:.1f} | {100 * t / tot:.1f} % |")
    except BrokenPipeError:
        pass
