#!/usr/bin/env python
"""Phase stamps of the general sweep (LSMC_TRACE=<file>, csrc/sweep.cu): per warp of CTA 0, averaged over tiles 2..62, in clocks.
F0 = wait for the tile + basis program, x~ = preconditioned regressors, fit = fitted values, bar1 = wait at the barrier before the
decision phase, D = own decision / Gram items, bar2 = wait at the end-of-tile barrier.  Usage: sweep_trace_report.py <file>"""
import sys

import numpy as np

t = np.fromfile(sys.argv[1], dtype=np.int64).reshape(16, 64, 8)
ok = [it for it in range(2, 63) if t[0, it, 0] and t[0, it + 1, 0]]
print("tiles used:", len(ok))
print("warp |    F0 |    x~ |   fit |  bar1 |     D |  bar2 | tile")
tot = []
for w in range(16):
    r = t[w][ok]
    nxt = t[w][[i + 1 for i in ok]]
    f0, xt, fit, bar1, d, bar2 = r[:, 1] - r[:, 0], r[:, 6] - r[:, 1], r[:, 2] - r[:, 6], r[:, 3] - r[:, 2], r[:, 4] - r[:, 3], r[:, 5] - r[:, 4]
    tile = nxt[:, 0] - r[:, 0]
    tot.append(tile.mean())
    print("%4d | %5.0f | %5.0f | %5.0f | %5.0f | %5.0f | %5.0f | %5.0f" % (w, f0.mean(), xt.mean(), fit.mean(), bar1.mean(), d.mean(), bar2.mean(), tile.mean()))
print("clocks per tile: %.0f" % np.mean(tot))
if t[13:, ok, 7].any():   # warps 13..15: next tile's rows landed (after the start of the decision phase) and its basis program done
    for w in (13, 14, 15):
        r = t[w][ok]
        print("warp %d: decision phase start -> next tile landed %.0f, basis program of the next tile %.0f" % (w, (r[:, 7] - r[:, 3]).mean(), (r[:, 4] - r[:, 7]).mean()))
