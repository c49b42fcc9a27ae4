"""SURVEY 8(f) row 4 on the GPU: point covariances (ModularSBASolver::std_step, modularsbasolver.cpp:952-1004) and the direct
reduced solve (the reference's default solver type ceres::SPARSE_SCHUR factorises S, modularsbasolver.h:672) -- both from a
hand-written dense FP64 Cholesky of the explicit reduced camera system (sba_kernels_dense.cu) -- against the oracle."""
import numpy as np
import pytest

from conftest import relerr
from steviapp_b200 import SBASolver, SBAError, synthetic

pytestmark = pytest.mark.gpu

FIXED_ITERS = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)


def gpu(sc, **opts):
    s = SBASolver(**opts)
    arrays = synthetic.load_into(s, sc)
    return s, arrays


@pytest.mark.parametrize("shape,kw", [("small", dict()), ("small", dict(mixed=True)), ("ladybug", dict())])
def test_direct_solve_lm_matches_oracle_cholesky(oracle, shape, kw):
    sc = synthetic.make_scene(shape, **kw)
    so = oracle.ExtProblem(sc).solve(max_iterations=5, num_threads=1, linear_solver=1, **FIXED_ITERS)
    s, _ = gpu(sc, linear_solver=1, operator_mode=3)
    out = s.solve(max_iterations=5, **FIXED_ITERS)
    assert out["pcg_iterations_total"] == 0
    for a, b in zip(out["cost_history"], so["cost_history"]):
        assert abs(a - b) <= 1e-8 * max(b, 1.0), (out["cost_history"], so["cost_history"])
    # and against the CG path of the engine itself (tight tolerance): the same steps
    s2, _ = gpu(sc, operator_mode=3, pcg_rel_tol=1e-13)
    out2 = s2.solve(max_iterations=5, **FIXED_ITERS)
    for a, b in zip(out["cost_history"], out2["cost_history"]):
        assert abs(a - b) <= 1e-8 * max(b, 1.0)


def test_direct_solve_with_free_intrinsics_and_pairs(oracle):
    from test_pairs_host import pair_scene
    sc = pair_scene(n_pairs=300)
    sc["intr_fixed_mask"] = np.array([0x10], np.uint8)
    intr = sc["intrinsics"].copy(); intr[0, 0] *= 1.002; intr[0, 3] += 0.002
    sc["intrinsics"] = intr
    sc["loss"] = "cauchy"
    p = oracle.ExtProblem(sc)
    so = p.solve(max_iterations=5, num_threads=1, linear_solver=1, **FIXED_ITERS)
    s, _ = gpu(sc, linear_solver=1)
    out = s.solve(max_iterations=5, **FIXED_ITERS)
    for a, b in zip(out["cost_history"], so["cost_history"]):
        assert abs(a - b) <= 1e-8 * max(b, 1.0), (out["cost_history"], so["cost_history"])
    assert relerr(s.intrinsics, p.intr.reshape(-1, 10)) < 1e-7


@pytest.mark.parametrize("case", ["small", "ladybug", "free_intrinsics", "heavy"])
def test_point_covariances(oracle, case):
    if case == "heavy":
        from test_gpu_parity import _add_heavy_points
        sc = _add_heavy_points(synthetic.make_scene(None, n_cams=70, n_pts=300, n_obs=1800, seed=5), [0, 7])
    else:
        sc = synthetic.make_scene("ladybug" if case == "ladybug" else "small")
    sc["pt_fixed"] = sc["pt_fixed"].copy(); sc["pt_fixed"][9] = 1
    if case == "free_intrinsics":
        sc["intr_fixed_mask"] = np.array([0x1c], np.uint8)      # f, pp free
    p = oracle.ExtProblem(sc)
    p.solve(max_iterations=4, num_threads=1, **FIXED_ITERS)
    co = p.point_covariance()
    s, _ = gpu(sc, operator_mode=3)
    s.solve(max_iterations=4, **FIXED_ITERS)
    c = s.point_covariances()
    assert np.abs(c[9]).max() == 0.0 and np.abs(co[9]).max() == 0.0
    scale = np.abs(co).max(axis=(1, 2), keepdims=True) + 1e-300
    assert np.abs((c - co) / scale).max() < 1e-6, np.abs((c - co) / scale).max()
    assert np.abs(c - np.transpose(c, (0, 2, 1))).max() <= 1e-9 * np.abs(c).max()
    free = np.flatnonzero(sc["pt_fixed"] == 0)
    assert (np.linalg.eigvalsh(c[free]) > 0).all()


def test_covariance_fails_loudly_without_a_gauge():
    """No fixed camera and no prior: J^T J is rank deficient -> "Failed to compute covariance matrix" (modularsbasolver.cpp:995-998)."""
    sc = synthetic.make_scene("small")
    sc["cam_fixed"] = np.zeros_like(sc["cam_fixed"])
    sc["prior_kind"] = np.zeros(0, np.int32); sc["prior_index"] = np.zeros(0, np.int32); sc["prior_A"] = np.zeros((0, 9)); sc["prior_b"] = np.zeros((0, 3))
    s, _ = gpu(sc, operator_mode=3)
    try:
        c = s.point_covariances()
    except SBAError as e:
        assert e.code == -5 and "covariance" in str(e)
    else:   # rounding can leave the 7 gauge directions barely positive: then the variances explode instead
        assert np.abs(c).max() > 1e6


def test_venice_scale_direct_solve_and_covariance():
    """Full BASELINE size: 10 752^2 dense factor.  The direct step equals the CG step (cost history 1e-8), the covariances are
    symmetric positive definite and of the size the pixel noise implies."""
    sc = synthetic.make_scene("venice")
    s, _ = gpu(sc)
    a = s.solve(max_iterations=2, **FIXED_ITERS)
    s.close()
    s, _ = gpu(sc, linear_solver=1)
    b = s.solve(max_iterations=2, **FIXED_ITERS)
    for x, y in zip(a["cost_history"], b["cost_history"]):
        assert abs(x - y) <= 1e-8 * y, (a["cost_history"], b["cost_history"])
    c = s.point_covariances()
    assert np.isfinite(c).all()
    assert np.abs(c - np.transpose(c, (0, 2, 1))).max() <= 1e-9 * np.abs(c).max()
    sig = np.sqrt(np.einsum("nii->ni", c))
    assert (sig > 0).all() and np.median(sig) < 0.05      # 0.5 px at f = 2000 and ~30 units range: millimetre-to-centimetre level
    print("direct solve ms per LM iteration:", b["ms_total"] / 2, "PCG path:", a["ms_total"] / 2, "median sigma:", np.median(sig))
