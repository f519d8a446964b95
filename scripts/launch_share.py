#!/usr/bin/env python
"""Aggregates an ncu launch list (`--metrics gpu__time_duration.sum --csv`) per kernel: launches, total and mean
duration, share of the captured window.  Usage: scripts/launch_share.py gpurun_out/launches_<tag>.csv"""
import csv
import re
import sys
from collections import OrderedDict


def main(path):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    ik, iv, ig, ib = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
    agg = OrderedDict()
    for r in rows[1:]:
        name = re.sub(r"\(.*$", "", r[ik]).replace("void ", "")
        a = agg.setdefault(name, [0, 0.0, r[ig], r[ib]])
        a[0] += 1
        a[1] += float(r[iv].replace(",", ""))
    total = sum(a[1] for a in agg.values())
    print("%-40s %8s %12s %12s %8s  %s" % ("kernel", "launches", "total ms", "mean us", "share", "grid x block"))
    for name, (n, ns, g, b) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-40s %8d %12.3f %12.2f %7.1f%%  %s x %s" % (name[:40], n, ns / 1e6, ns / n / 1e3, 100 * ns / total, g, b))
    print("%-40s %8d %12.3f" % ("captured window", sum(a[0] for a in agg.values()), total / 1e6))


if __name__ == "__main__":
    main(sys.argv[1])
