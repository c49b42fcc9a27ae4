"""Preconditioner study on the explicit reduced camera system of a real LM iteration (numpy/scipy on the host):
PCG iteration counts of the 6x6 block-Jacobi (SCHUR_JACOBI) vs larger diagonal clusters.  Usage: precond_study.py [shape] [lm_iters]"""
import json
import os
import sys
import time

import numpy as np
import scipy.linalg as sl

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
lm = int(sys.argv[2]) if len(sys.argv) > 2 else 3
sc = synthetic.make_scene(shape)
fixed = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
with SBASolver(pcg_check_every=8, operator_mode=3, **fixed) as s:
    synthetic.load_into(s, sc)
    r = s.solve(max_iterations=lm)
    S, b = s.explicit_dense()
print("engine pcg its", r["pcg_iters_history"][:lm + 1], "n", S.shape[0], flush=True)
free = np.repeat(~sc["cam_fixed"].astype(bool), 6)
S = S[np.ix_(free, free)]; b = b[free]
n = S.shape[0]


def pcg(apply_minv, tol=1e-10, maxit=1000):
    x = np.zeros(n); rr = b.copy(); z = apply_minv(rr); p = z.copy(); rz = rr @ z; rz0 = rz
    for it in range(1, maxit + 1):
        q = S @ p
        a = rz / (p @ q)
        x += a * p; rr -= a * q
        z = apply_minv(rr); rzn = rr @ z
        if not rzn > tol * tol * rz0:
            return it
        p = z + (rzn / rz) * p; rz = rzn
    return maxit


def cluster_jacobi(cams_per_cluster, shift=0):
    w = 6 * cams_per_cluster
    blocks = []
    start = 0
    edges = list(range(6 * shift, n, w)) if shift else list(range(0, n, w))
    if edges[0] != 0:
        edges = [0] + edges
    edges.append(n)
    for a, e in zip(edges[:-1], edges[1:]):
        blocks.append((a, e, sl.cho_factor(S[a:e, a:e])))

    def apply(r):
        z = np.empty_like(r)
        for a, e, c in blocks:
            z[a:e] = sl.cho_solve(c, r[a:e])
        return z
    return apply


out = {}
for cpc in (1, 4, 16, 32, 64, 128):
    t0 = time.time()
    its = pcg(cluster_jacobi(cpc))
    out[cpc] = its
    print(f"cluster of {cpc:4d} cameras: {its} PCG iterations ({time.time() - t0:.1f} s)", flush=True)
json.dump(out, open(os.path.join("gpurun_out", f"precond_study_{shape}.json"), "w"))
