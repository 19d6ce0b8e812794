"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name (shares, not absolutes)."""
import collections
import csv
import sys

for path in sys.argv[1:]:
    rows = list(csv.reader(open(path)))
    hdr = [r for r in rows if "Kernel Name" in r][0]
    k, v = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg, cnt = collections.Counter(), collections.Counter()
    for r in rows:
        if len(r) == len(hdr) and r[v].replace(".", "").isdigit():
            n = r[k].split("(")[0].replace("void ", "").replace("<unnamed>::", "")[-70:]
            agg[n] += float(r[v])
            cnt[n] += 1
    tot = sum(agg.values())
    print(f"## {path}: {sum(cnt.values())} launches, {tot / 1e3:.1f} us in kernels")
    print("| kernel | launches | us | share |\n|---|---|---|---|")
    for n, t in agg.most_common(16):
        print(f"| `{n}` | {cnt[n]} | {t / 1e3:.1f} | {100 * t / tot:.1f} % |")
