#!/usr/bin/env python3
"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel:
launches, total ms, share.  Usage: launch_summary.py raw.csv > summary.csv"""
import collections
import csv
import re
import sys


def main():
    lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
    r = csv.reader(lines)
    hdr = next(r)
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "second": 1e3}
    agg = collections.OrderedDict()
    order = []
    for row in r:
        name = row[ki]
        ms = float(row[vi].replace(",", "")) * scale.get(row[ui], 1.0)
        order.append((name, ms))
        a = agg.setdefault(name, [0, 0.0, 1e30, 0.0])
        a[0] += 1
        a[1] += ms
        a[2] = min(a[2], ms)
        a[3] = max(a[3], ms)
    total = sum(a[1] for a in agg.values())
    w = csv.writer(sys.stdout)
    w.writerow(["kernel", "launches", "total_ms", "share_pct", "min_ms", "max_ms"])
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        w.writerow([re.sub(r"\s+", " ", k), a[0], "%.4f" % a[1], "%.2f" % (100 * a[1] / total),
                    "%.4f" % a[2], "%.4f" % a[3]])
    return 0


if __name__ == "__main__":
    sys.exit(main())
