"""The product's closed-form residual/Jacobian math (steviapp_b200/csrc/sba_math.cuh, compiled for the
host by tests/hostmath) against the oracle's forward-mode autodiff of the reference functor.
Gate: 1e-12 relative (north star).  CPU only."""
import numpy as np
import pytest

from conftest import ptr, relerr
from steviapp_b200 import synthetic

TOL = 1e-12


def hm_eval(hostmath, sc):
    n = sc["obs_cam"].size
    res = np.empty((n, 2)); jc = np.empty((n, 2, 6)); jp = np.empty((n, 2, 3)); failed = np.zeros(n, np.uint8)
    info = None if sc["obs_info"] is None else np.ascontiguousarray(sc["obs_info"], dtype=np.float64)
    bad = hostmath.hm_evaluate(n, ptr(sc["obs_cam"]), ptr(sc["obs_pt"]), ptr(sc["raxis"]), ptr(sc["t"]),
                               ptr(sc["cost_kind"]), ptr(sc["cam_model"]), ptr(sc["intrinsics"]), ptr(sc["pts"]),
                               ptr(sc["obs_uv"]), ptr(info), ptr(res), ptr(jc), ptr(jp), ptr(failed))
    return res, jc, jp, failed, bad


@pytest.mark.parametrize("kw", [dict(), dict(spherical=True), dict(mixed=True)])
@pytest.mark.parametrize("with_info", [False, True])
def test_closed_form_matches_autodiff(hostmath, oracle, kw, with_info):
    sc = synthetic.make_scene("small", **kw)
    if with_info:
        rng = np.random.default_rng(5)
        sc["obs_info"] = rng.normal(size=(sc["n_obs"], 2, 2)) + 2 * np.eye(2)
    o = oracle.Problem(sc).evaluate()
    res, jc, jp, failed, bad = hm_eval(hostmath, sc)
    assert bad == 0 and o["ok"]
    assert relerr(res, o["residuals"]) < TOL
    assert relerr(jc, o["jac_cam"]) < TOL
    assert relerr(jp, o["jac_pt"]) < TOL


@pytest.mark.parametrize("scale", [1e-9, 1e-6, 1e-4, 1e-3, 1e-2, 0.3, 0.5, 3.0, np.pi - 1e-3, np.pi - 1e-6])
def test_rotation_branch_regions(hostmath, oracle, scale):
    """|r| swept through the small-angle / series / closed-form branches of so3_coeffs and up to pi."""
    sc = synthetic.make_scene("tiny")
    rng = np.random.default_rng(11)
    n = sc["n_cams"]
    ax = rng.normal(size=(n, 3)); ax /= np.linalg.norm(ax, axis=1)[:, None]
    sc["raxis"] = np.ascontiguousarray(ax * scale)
    sc["cost_kind"] = np.ones(n, np.uint8)   # spherical: no cheirality constraint for arbitrary poses
    o = oracle.Problem(sc).evaluate()
    res, jc, jp, failed, bad = hm_eval(hostmath, sc)
    assert relerr(res, o["residuals"]) < TOL
    assert relerr(jc, o["jac_cam"]) < 1e-11 if scale > 3.1 else relerr(jc, o["jac_cam"]) < TOL
    assert relerr(jp, o["jac_pt"]) < TOL


def test_cheirality_flag(hostmath, oracle):
    sc = synthetic.make_scene("tiny")
    sc["pts"] = sc["pts"].copy()
    sc["pts"][0] = sc["t"][sc["obs_cam"][sc["obs_pt"] == 0][0]] * 3.0   # behind its first camera
    o = oracle.Problem(sc).evaluate()
    res, jc, jp, failed, bad = hm_eval(hostmath, sc)
    assert not o["ok"] and bad > 0
    assert np.array_equal(failed, o["failed"])


def test_inv6_and_inv3(hostmath):
    rng = np.random.default_rng(2)
    for _ in range(20):
        B = rng.normal(size=(6, 6)); A = B @ B.T + 0.1 * np.eye(6)
        Ai = np.empty((6, 6))
        assert hostmath.hm_inv6(ptr(np.ascontiguousarray(A)), ptr(Ai)) == 0
        assert np.abs(Ai @ A - np.eye(6)).max() < 1e-9
        B = rng.normal(size=(3, 3)); H = B @ B.T + 0.1 * np.eye(3)
        h6 = np.array([H[0, 0], H[0, 1], H[0, 2], H[1, 1], H[1, 2], H[2, 2]]); hi = np.empty(6)
        assert hostmath.hm_inv3(ptr(h6), ptr(hi)) == 0
        Hi = np.array([[hi[0], hi[1], hi[2]], [hi[1], hi[3], hi[4]], [hi[2], hi[4], hi[5]]])
        assert np.abs(Hi @ H - np.eye(3)).max() < 1e-10
    assert hostmath.hm_inv6(ptr(np.zeros((6, 6))), ptr(np.empty((6, 6)))) == 1


@pytest.mark.parametrize("kw", [dict(), dict(spherical=True), dict(mixed=True)])
@pytest.mark.parametrize("with_info", [False, True])
def test_intrinsics_jacobian_closed_form_matches_autodiff(hostmath, oracle, kw, with_info):
    """SURVEY 8(f) row 1, first half: d res / d (f, pp, k, t, B) in closed form (obs_intr_jac) against the duals of the reference
    functor over its five intrinsic blocks; the B columns are identically zero (pinholecameraprojectormodule.h:46-47)."""
    sc = synthetic.make_scene("small", **kw)
    if with_info:
        rng = np.random.default_rng(5)
        sc["obs_info"] = rng.normal(size=(sc["n_obs"], 2, 2)) + 2 * np.eye(2)
    jo = oracle.Problem(sc).intrinsics_jac()
    n = sc["obs_cam"].size
    j = np.empty((n, 2, 10))
    info = None if sc["obs_info"] is None else np.ascontiguousarray(sc["obs_info"], dtype=np.float64)
    bad = hostmath.hm_intrinsics_jac(n, ptr(sc["obs_cam"]), ptr(sc["obs_pt"]), ptr(sc["raxis"]), ptr(sc["t"]), ptr(sc["cost_kind"]),
                                     ptr(sc["cam_model"]), ptr(sc["intrinsics"]), ptr(sc["pts"]), ptr(sc["obs_uv"]), ptr(info), ptr(j))
    assert bad == 0
    assert np.abs(jo[:, :, 8:]).max() == 0.0 and np.abs(j[:, :, 8:]).max() == 0.0
    # column-wise relative error: the columns span ~8 orders of magnitude (d/dpp ~ 1, d/dk3 ~ f r^6)
    for c in range(8):
        assert relerr(j[:, :, c], jo[:, :, c]) < TOL, c
