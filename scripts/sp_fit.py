"""Least-squares fit of the per-CTA mat-vec time of the block-sparse persistent PCG kernel on the composition of the runs
(SBA_SP_DUMP=1 SBA_PS_CLOCK=2 output of scripts/prof_target.py).  Usage: sp_fit.py <log>"""
import re
import sys

import numpy as np

txt = open(sys.argv[1]).read()
comp = {}
for m in re.finditer(r"\[sp_dump\] (\d+) (\d+) (\d+) (\d+) (\d+) (\d+) (\d+)", txt):
    v = [int(x) for x in m.groups()]
    comp[v[0]] = v[1:]
times = {}
line = [l for l in txt.splitlines() if "per-CTA mat-vec us" in l][-1]
for m in re.finditer(r"(\d+):(\d+)=([\d.]+)", line):
    times[int(m.group(1))] = float(m.group(3))
ids = sorted(comp)
X = np.array([comp[i] for i in ids], float)
t = np.array([times[i] for i in ids])
names = ["bytes", "panels", "chunks", "segs", "loops", "wmax"]
print("time us: mean %.2f min %.2f max %.2f" % (t.mean(), t.min(), t.max()))
for k, nme in enumerate(names):
    print("%-7s mean %10.1f min %10.1f max %10.1f corr with time %.3f" % (nme, X[:, k].mean(), X[:, k].min(), X[:, k].max(), np.corrcoef(X[:, k], t)[0, 1]))
for cols in ([0, 1, 3], [0, 1, 2, 3], [0, 3, 4], [0, 3, 5], [0, 1, 3, 5], [0, 1, 2, 3, 4, 5]):
    A = X[:, cols]
    c, res, *_ = np.linalg.lstsq(A, t, rcond=None)
    pred = A @ c
    print([names[k] for k in cols], "coef (us per unit)", c, "rms resid %.2f us" % np.sqrt(np.mean((pred - t) ** 2)),
          "byte equivalents:", (c / c[0]).round(0))
