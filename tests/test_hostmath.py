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


@pytest.mark.parametrize("spherical", [0, 1])
@pytest.mark.parametrize("with_info", [False, True])
def test_constant_sensor_offset_matches_autodiff(hostmath, oracle, spherical, with_info):
    """SURVEY 8(f) row 3, math half: the closed form of the PoseTransform<UVCost, SourceToInitial> cost (obs_eval_offset)
    against the oracle's duals over the reference functor, 1e-12.  Random body poses, offsets of up to ~1 rad / 0.5 units,
    points in front of the composed sensor pose; identity offset reproduces the plain cost."""
    rng = np.random.default_rng(17 + spherical)
    n = 400
    intr = np.array([2000.0, 2000.0, 1500.0, -0.05, 0.01, 0.001, 1e-4, -1e-4, 0.0, 0.0])

    def unit(k):
        v = rng.normal(size=(k, 3))
        return v / np.linalg.norm(v, axis=1)[:, None]

    r = unit(n) * rng.uniform(0.05, 3.0, (n, 1))
    t = rng.normal(0, 2.0, (n, 3))
    r_off = unit(n) * rng.uniform(0.0, 1.0, (n, 1))
    t_off = rng.normal(0, 0.5, (n, 3))
    r_off[:20] = 0.0; t_off[:20] = 0.0                     # identity offsets
    # sensor pose = pose * offset; points in front of it, pixels = exact projection + noise
    Rb, Ro = synthetic.axis_angle_to_rotmat(r), synthetic.axis_angle_to_rotmat(r_off)
    Rs = Rb @ Ro
    ts = np.einsum("nij,nj->ni", Rb, t_off) + t
    Ps = np.column_stack([rng.uniform(-4, 4, n), rng.uniform(-3, 3, n), rng.uniform(5, 40, n)])
    pts = np.einsum("nij,nj->ni", Rs, Ps) + ts
    uv = synthetic.uv_from_normalized(Ps[:, :2] / Ps[:, 2:3], intr) + rng.normal(0, 0.5, (n, 2))
    uv = np.ascontiguousarray(uv, dtype=np.float64)
    info = np.ascontiguousarray(rng.normal(size=(n, 2, 2)) + 2 * np.eye(2)) if with_info else None
    off_rt = np.ascontiguousarray(np.concatenate(
        [np.transpose(Ro, (0, 2, 1)).reshape(n, 9), t_off], axis=1))

    lib = oracle.load()
    ro = np.empty((n, 2)); jco = np.empty((n, 2, 6)); jpo = np.empty((n, 2, 3)); fo = np.zeros(n, np.uint8)
    bad_o = lib.oracle_eval_offset(n, ptr(r), ptr(t), ptr(r_off), ptr(t_off), ptr(pts), ptr(intr), ptr(uv), ptr(info), spherical,
                                   ptr(ro), ptr(jco), ptr(jpo), ptr(fo))
    rh = np.empty((n, 2)); jch = np.empty((n, 2, 6)); jph = np.empty((n, 2, 3)); fh = np.zeros(n, np.uint8)
    bad_h = hostmath.hm_eval_offset(n, ptr(r), ptr(t), ptr(off_rt), ptr(pts), ptr(intr), ptr(uv), ptr(info), spherical,
                                    ptr(rh), ptr(jch), ptr(jph), ptr(fh))
    assert bad_o == 0 and bad_h == 0
    assert np.abs(ro).max() < 50.0                          # the samples are what they claim: near-exact projections
    # Jacobians at the 1e-12 gate; the residual is a difference of ~2000-pixel quantities formed through TWO rotations on
    # both sides (one ulp of 2000 is 4.5e-13), so its gate here is 2e-12
    assert relerr(rh, ro) < 2 * TOL and relerr(jch, jco) < TOL and relerr(jph, jpo) < TOL
    # identity offset = the plain cost (same closed form as obs_eval, same oracle functor as uv_cost)
    sc = dict(obs_cam=np.arange(20, dtype=np.int32), obs_pt=np.arange(20, dtype=np.int32), raxis=r[:20].copy(), t=t[:20].copy(),
              cost_kind=np.full(20, spherical, np.uint8), cam_model=np.zeros(20, np.int32), intrinsics=intr[None, :].copy(),
              pts=pts[:20].copy(), obs_uv=uv[:20].copy(), obs_info=None if info is None else info[:20].copy())
    res, jc, jp, failed, bad = hm_eval(hostmath, sc)
    assert bad == 0 and relerr(rh[:20], res) < 1e-13 and relerr(jch[:20], jc) < 1e-13 and relerr(jph[:20], jp) < 1e-13


def test_constant_sensor_offset_cheirality(hostmath, oracle):
    """Plane cost: the functor fails when the point is behind the COMPOSED sensor pose (offset turned by pi about y)."""
    intr = np.array([35.0, 320.0, 240.0, 0, 0, 0, 0, 0, 0, 0])
    r = np.zeros((1, 3)); t = np.zeros((1, 3)); pts = np.array([[1.0, 2.0, 10.0]]); uv = np.array([[323.5, 247.0]])
    r_off = np.array([[0.0, np.pi, 0.0]]); t_off = np.zeros((1, 3))
    off_rt = np.ascontiguousarray(np.concatenate([synthetic.axis_angle_to_rotmat(r_off)[0].T.reshape(1, 9), t_off], axis=1))
    out = [np.empty((1, 2)), np.empty((1, 2, 6)), np.empty((1, 2, 3)), np.zeros(1, np.uint8)]
    assert oracle.load().oracle_eval_offset(1, ptr(r), ptr(t), ptr(r_off), ptr(t_off), ptr(pts), ptr(intr), ptr(uv), None, 0,
                                            *[ptr(a) for a in out]) == 1
    assert hostmath.hm_eval_offset(1, ptr(r), ptr(t), ptr(off_rt), ptr(pts), ptr(intr), ptr(uv), None, 0, *[ptr(a) for a in out]) == 1


@pytest.mark.parametrize("spherical", [0, 1])
@pytest.mark.parametrize("with_info", [False, True])
def test_lever_arm_behind_offset_matches_autodiff(hostmath, oracle, spherical, with_info):
    """SURVEY 8(f) row 3, math half: PoseTransform<LeverArm<UVCost>> -- free boresight and lever arm behind a constant offset --
    as a closed form (obs_eval_leverarm) against the oracle's duals over the 15 free scalars; with boresight = lever arm = 0
    it reproduces the constant-offset cost."""
    rng = np.random.default_rng(29 + spherical)
    n = 400
    intr = np.array([2000.0, 2000.0, 1500.0, -0.05, 0.01, 0.001, 1e-4, -1e-4, 0.0, 0.0])

    def unit(k):
        v = rng.normal(size=(k, 3))
        return v / np.linalg.norm(v, axis=1)[:, None]

    r = unit(n) * rng.uniform(0.05, 3.0, (n, 1)); t = rng.normal(0, 2.0, (n, 3))
    r_off = unit(n) * rng.uniform(0.0, 1.0, (n, 1)); t_off = rng.normal(0, 0.5, (n, 3))
    b = unit(n) * rng.uniform(0.0, 0.5, (n, 1)); l = rng.normal(0, 0.3, (n, 3))
    b[:20] = 0.0; l[:20] = 0.0
    Rb, Ro, Rl = (synthetic.axis_angle_to_rotmat(a) for a in (r, r_off, b))
    Rs = Rb @ Ro @ Rl
    ts = np.einsum("nij,nj->ni", Rb, np.einsum("nij,nj->ni", Ro, l) + t_off) + t
    Ps = np.column_stack([rng.uniform(-4, 4, n), rng.uniform(-3, 3, n), rng.uniform(5, 40, n)])
    pts = np.einsum("nij,nj->ni", Rs, Ps) + ts
    uv = np.ascontiguousarray(synthetic.uv_from_normalized(Ps[:, :2] / Ps[:, 2:3], intr) + rng.normal(0, 0.5, (n, 2)))
    info = np.ascontiguousarray(rng.normal(size=(n, 2, 2)) + 2 * np.eye(2)) if with_info else None
    off_rt = np.ascontiguousarray(np.concatenate([np.transpose(Ro, (0, 2, 1)).reshape(n, 9), t_off], axis=1))

    def outs():
        return [np.empty((n, 2)), np.empty((n, 2, 6)), np.empty((n, 2, 6)), np.empty((n, 2, 3)), np.zeros(n, np.uint8)]

    o = outs()
    assert oracle.load().oracle_eval_leverarm(n, ptr(r), ptr(t), ptr(b), ptr(l), ptr(r_off), ptr(t_off), ptr(pts), ptr(intr), ptr(uv),
                                              ptr(info), spherical, *[ptr(a) for a in o]) == 0
    h = outs()
    assert hostmath.hm_eval_leverarm(n, ptr(r), ptr(t), ptr(b), ptr(l), ptr(off_rt), ptr(pts), ptr(intr), ptr(uv), ptr(info), spherical,
                                     *[ptr(a) for a in h]) == 0
    assert np.abs(o[0]).max() < 50.0
    # residual gate 3e-12: a difference of ~2000-pixel quantities behind THREE rotations (see the offset test); Jacobians 1e-12
    assert relerr(h[0], o[0]) < 3 * TOL
    assert relerr(h[1], o[1]) < TOL and relerr(h[2], o[2]) < TOL and relerr(h[3], o[3]) < TOL
    # boresight = lever arm = 0: the constant-offset closed form
    k = 20
    rh = np.empty((k, 2)); jch = np.empty((k, 2, 6)); jph = np.empty((k, 2, 3)); fh = np.zeros(k, np.uint8)
    assert hostmath.hm_eval_offset(k, ptr(r[:k].copy()), ptr(t[:k].copy()), ptr(off_rt[:k].copy()), ptr(pts[:k].copy()), ptr(intr),
                                   ptr(uv[:k].copy()), ptr(None if info is None else info[:k].copy()), spherical,
                                   ptr(rh), ptr(jch), ptr(jph), ptr(fh)) == 0
    assert relerr(h[0][:k], rh) < 1e-13 and relerr(h[1][:k], jch) < 1e-13 and relerr(h[3][:k], jph) < 1e-13
