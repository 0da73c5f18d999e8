#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum --csv): python tools/launch_summary.py file.csv [...]"""
import collections
import csv
import sys


def main():
    for f in sys.argv[1:]:
        rows = [r for r in csv.reader(open(f)) if len(r) > 5]
        hdr = None
        agg = collections.defaultdict(lambda: [0, 0.0])
        for r in rows:
            if r[0] == "ID":
                hdr = r
                continue
            if hdr is None:
                continue
            name = r[hdr.index("Kernel Name")].split("(")[0].replace("void ", "").replace("wf::", "")[:60]
            v = float(r[hdr.index("Metric Value")].replace(",", ""))
            u = r[hdr.index("Metric Unit")]
            v = v / 1000 if u == "ns" else v * 1000 if u == "ms" else v
            agg[name][0] += 1
            agg[name][1] += v
        tot = sum(v[1] for v in agg.values())
        print(f, "serialised total us", round(tot, 1))
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            print(f"  {k:60s} n={v[0]:4d} avg={v[1] / v[0]:9.2f} us  share={v[1] / tot:.3f}")


if __name__ == "__main__":
    main()
