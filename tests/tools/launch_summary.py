"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) by kernel name.

    python tests/tools/launch_summary.py profiles/r01b_launches_bench_nograph.csv [--skip N] [--take M]
"""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    rows = list(csv.reader(open(path, errors="replace")))
    hdr = None
    items = []
    for r in rows:
        if "Kernel Name" in r:
            hdr = r
            continue
        if hdr is None or len(r) != len(hdr):
            continue
        d = dict(zip(hdr, r))
        try:
            v = float(d["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        name = re.sub(r"\(CUtensorMap.*|\(.*", "", d["Kernel Name"]).replace("void ", "")
        items.append((name, v, d["Grid Size"]))
    agg = collections.OrderedDict()
    for n, v, g in items:
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    for n, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{v / 1e3:10.1f} us {c:5d}x {v / c / 1e3:8.1f} us/launch {v / tot * 100:5.1f}%  {n}")
    print(f"{tot / 1e3:10.1f} us total over {len(items)} launches")


if __name__ == "__main__":
    main()
