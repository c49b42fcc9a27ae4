"""Preconditioner study for the shared-block (free intrinsics) reduced system (numpy on the host, S from the engine's hook):
PCG iterations with (a) the 6x6 block-Jacobi the engine and the oracle use, (b) one joint block for the pseudo-cameras,
(c) the bordered preconditioner M = [D B; B^T Sgg] (D = 6x6 camera blocks, B = border columns) applied exactly."""
import os
import sys

import numpy as np
import scipy.linalg as sl

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

sc = synthetic.make_scene(sys.argv[1] if len(sys.argv) > 1 else "venice")
sc["intr_fixed_mask"] = np.array([0x10], np.uint8)
lm = int(sys.argv[2]) if len(sys.argv) > 2 else 2
fixed = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
with SBASolver(pcg_check_every=8, operator_mode=3, **fixed) as s:
    synthetic.load_into(s, sc)
    r = s.solve(max_iterations=lm)
    S, b = s.explicit_dense()
    info = s.shared_info()
nc, next_ = sc["n_cams"], info["n_ext"]
print("engine pcg its", r["pcg_iters_history"][:lm + 1], "n", S.shape, "cams", nc, "ext", next_, flush=True)
live = np.abs(np.diag(S)) > 0
live &= ~((np.abs(S - np.diag(np.diag(S))).sum(axis=1) == 0) & (b == 0))        # dead columns: unit diagonal, zero rhs
S = S[np.ix_(live, live)]; b = b[live]
n = S.shape[0]
idx = np.flatnonzero(live)
is_g = idx >= 6 * nc
ng = int(is_g.sum())
print("live columns", n, "of which shared", ng, flush=True)


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


print("symmetry defect", np.abs(S - S.T).max(), "min diag", np.diag(S).min(), flush=True)
gg = S[np.ix_(np.flatnonzero(is_g), np.flatnonzero(is_g))]
print("shared block diag", np.diag(gg), "eig", np.linalg.eigvalsh(gg), flush=True)
cam_cols = np.flatnonzero(~is_g); g_cols = np.flatnonzero(is_g)
cam_of = idx[cam_cols] // 6
blocks = [cam_cols[cam_of == c] for c in np.unique(cam_of)]
chol = [sl.cho_factor(S[np.ix_(bk, bk)]) for bk in blocks]
pc_of = idx[g_cols] // 6
gblocks = [g_cols[pc_of == c] for c in np.unique(pc_of)]
gchol = [sl.cho_factor(S[np.ix_(bk, bk)]) for bk in gblocks]
Sgg = sl.cho_factor(S[np.ix_(g_cols, g_cols)])
B = S[np.ix_(cam_cols, g_cols)]


def dinv(rc):
    z = np.empty_like(rc)
    for bk, c in zip(blocks, chol):
        loc = np.searchsorted(cam_cols, bk)
        z[loc] = sl.cho_solve(c, rc[loc])
    return z


def jacobi66(r):
    z = np.empty_like(r)
    z[cam_cols] = dinv(r[cam_cols])
    for bk, c in zip(gblocks, gchol):
        z[bk] = sl.cho_solve(c, r[bk])
    return z


def joint_g(r):
    z = np.empty_like(r)
    z[cam_cols] = dinv(r[cam_cols])
    z[g_cols] = sl.cho_solve(Sgg, r[g_cols])
    return z


DinvB = np.column_stack([dinv(B[:, k]) for k in range(ng)])
Sg_mat = S[np.ix_(g_cols, g_cols)] - B.T @ DinvB
print("eigenvalues of Sgg - B^T D^-1 B:", np.linalg.eigvalsh(Sg_mat), flush=True)
try:
    Sg = sl.cho_factor(Sg_mat)
except np.linalg.LinAlgError:
    Sg = None
    print("bordered preconditioner [D B; B^T Sgg] is NOT positive definite (D drops the camera-camera coupling): unusable for CG", flush=True)


def bordered(r):
    z = np.empty_like(r)
    yc = dinv(r[cam_cols])
    zg = sl.cho_solve(Sg, r[g_cols] - B.T @ yc)
    z[g_cols] = zg
    z[cam_cols] = yc - DinvB @ zg
    return z


for name, f in (("6x6 block-Jacobi (engine / oracle)", jacobi66), ("cameras 6x6 + one joint block for the shared columns", joint_g),
                ("bordered: exact [D B; B^T Sgg]", bordered)):
    if f is bordered and Sg is None:
        continue
    print(f"{name}: {pcg(f)} PCG iterations", flush=True)
