#!/usr/bin/env python
"""Development: prints the pipeline timeline recorded by LSMC_TRACE (per-warp phase timestamps of CTA 0 of one sweep)."""
import sys
import numpy as np
a = np.fromfile(sys.argv[1], dtype=np.int64).reshape(16, 64, 8)
entry, waited = int(a[0, 63, 6]), int(a[0, 63, 7])
a[0, 63, 6] = a[0, 63, 7] = 0
if entry > 0:
    body = a[:13, :, :4]
    print("kernel entry -> dependency wait over: %d clk; -> first decision tile starts: %d clk; -> last decision tile ends: %d clk" % (
        waited - entry, body[:, 0, 2][body[:, 0, 2] > 0].min() - entry, body[:, :, 3].max() - entry))
t0 = a[a > 0].min()
rel = np.where(a > 0, a - t0, -1)
ntile = int((rel[0, :, 0] >= 0).sum())
print("tiles recorded:", ntile)
print("decision warps: [wait cready] [wait vfull] [compute] per tile (clk); aux: [wait dfree] [wait raw] [products+bars] [dmma]")
for it in range(2, min(ntile, 14)):
    row = []
    for w in (0, 1, 4, 12):
        r = rel[w, it]
        row.append("w%-2d st %6d wc %5d wv %4d cmp %5d" % (w, r[0], r[1] - r[0], r[2] - r[1], r[3] - r[2]))
    print("it %2d | " % it + " | ".join(row))
    row = []
    for w in (13, 14, 15):
        r = rel[w, it]
        row.append("a%-2d st %6d wd %5d wr %4d b1 %5d prod %5d b2 %5d mm %5d end %6d" % (w, r[0], r[1] - r[0], r[2] - r[1], r[5] - r[2], r[6] - r[5], r[3] - r[6], r[4] - r[3], r[4]))
    print("      | " + " | ".join(row))
d = rel[:13, 3:ntile - 1]
print("decision compute clk per tile: mean %.0f  (per warp: %s)" % ((d[:, :, 3] - d[:, :, 2]).mean(), np.round((d[:, :, 3] - d[:, :, 2]).mean(axis=1)).astype(int)))
print("decision wait-cready clk per tile: mean %.0f" % (d[:, :, 1] - d[:, :, 0]).mean())
x = rel[13:16, 3:ntile - 1]
print("aux: wait dfree %.0f, wait raw %.0f, products %.0f, dmma %s" % ((x[:, :, 1] - x[:, :, 0]).mean(), (x[:, :, 2] - x[:, :, 1]).mean(), (x[:, :, 3] - x[:, :, 2]).mean(), np.round((x[:, :, 4] - x[:, :, 3]).mean(axis=1)).astype(int)))
per = np.diff(rel[0, 3:ntile - 1, 0]).mean()
print("tile period (warp 0): %.0f clk" % per)
