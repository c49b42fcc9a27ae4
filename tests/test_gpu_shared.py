"""SURVEY 8(f) rows 1 and 3 on the GPU against the extended oracle (oracle/sba_oracle_ext.cpp): free intrinsic blocks
(PinholeCamProjModule::init, sbamodules/pinholecameraprojectormodule.cpp:302-533), constant sensor offsets and free
boresight / lever-arm blocks (addProjectionCostFunctionImpl, :58-214) INSIDE the LM -- the reduced camera system with its
pseudo-cameras, the cost history of a fixed number of LM iterations (1e-8), the PCG iteration counts and the updated blocks."""
import numpy as np
import pytest

from conftest import relerr
from steviapp_b200 import SBASolver, SBAError, synthetic

pytestmark = pytest.mark.gpu

TOL_COST = 1e-8
FIXED_ITERS = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)


def perturbed_intrinsics(sc, rng, n_models=1):
    intr = np.tile(sc["intrinsics"][0], (n_models, 1))
    intr[:, 0] *= 1.0 + rng.normal(0, 0.004, n_models)
    intr[:, 1:3] += rng.normal(0, 3.0, (n_models, 2))
    intr[:, 3] += rng.normal(0, 0.004, n_models)
    intr[:, 6:8] += rng.normal(0, 2e-5, (n_models, 2))
    return intr


def add_rig(sc, rng, offsets=True, n_mounts=0, mount_noise=(0.004, 0.02)):
    """Re-expresses the scene's cameras as BODY poses behind a constant sensor offset and / or a mounting (boresight, lever
    arm): sensor2world = (pose * offset) * (b, l) (posedecoratorfunctors.h:208,628-700), so the pixels stay consistent."""
    sc = dict(sc)
    nc = sc["n_cams"]
    off = np.zeros((nc, 6))
    if offsets:
        off[:, :3] = rng.normal(0, 0.2, (nc, 3)); off[:, 3:] = rng.normal(0, 0.3, (nc, 3))
        off[::5] = 0.0                                  # some cameras keep the identity offset (the reference's first branch)
    mount_b = rng.normal(0, 0.15, (n_mounts, 3)); mount_l = rng.normal(0, 0.25, (n_mounts, 3))
    cam_mount = np.full(nc, -1, dtype=np.int32)
    if n_mounts:
        cam_mount = (np.arange(nc) % (n_mounts + 1) - 1).astype(np.int32)   # every (n_mounts + 1)-th camera has no mounting
    Rs = synthetic.axis_angle_to_rotmat(sc["raxis_true"])
    Roff = synthetic.axis_angle_to_rotmat(off[:, :3])
    Rb = np.tile(np.eye(3), (nc, 1, 1)); l = np.zeros((nc, 3))
    has = cam_mount >= 0
    if n_mounts:
        Rb[has] = synthetic.axis_angle_to_rotmat(mount_b[cam_mount[has]]); l[has] = mount_l[cam_mount[has]]
    M = Roff @ Rb                                        # sensor -> body rotation
    R = Rs @ np.transpose(M, (0, 2, 1))
    t = sc["t_true"] - np.einsum("nij,nj->ni", R, off[:, 3:] + np.einsum("nij,nj->ni", Roff, l))
    r, th = synthetic.rotmat_to_axis_angle(R)
    assert ((th > 0.02) & (th < np.pi - 0.02)).all()
    sc["raxis_true"], sc["t_true"] = r, t
    sc["raxis"] = np.ascontiguousarray(r + rng.normal(0, 0.01, r.shape)); sc["t"] = np.ascontiguousarray(t + rng.normal(0, 0.05, t.shape))
    sc["raxis"][0], sc["t"][0] = r[0], t[0]
    if offsets:
        sc["cam_offset"] = off
    if n_mounts:
        sc["cam_mount"] = cam_mount
        sc["mount_b_true"], sc["mount_l_true"] = mount_b, mount_l
        sc["mount_b"] = mount_b + rng.normal(0, mount_noise[0], mount_b.shape)
        sc["mount_l"] = mount_l + rng.normal(0, mount_noise[1], mount_l.shape)
    return sc


def with_pose_priors(sc, sigma_r=0.002, sigma_t=0.02):
    """Trajectory-style priors on every body pose (imagealignementsbamodule.cpp:256-286): without them a free lever arm is a
    gauge freedom of the free body poses."""
    sc = dict(sc)
    nc = sc["n_cams"]
    kind = np.concatenate([np.full(nc, 1, np.int32), np.full(nc, 2, np.int32)])
    index = np.concatenate([np.arange(nc, dtype=np.int32)] * 2)
    A = np.concatenate([np.tile((np.eye(3) / sigma_r).ravel(), (nc, 1)), np.tile((np.eye(3) / sigma_t).ravel(), (nc, 1))])
    b = np.concatenate([sc["raxis_true"], sc["t_true"]])
    sc["prior_kind"] = np.concatenate([sc["prior_kind"], kind]); sc["prior_index"] = np.concatenate([sc["prior_index"], index])
    sc["prior_A"] = np.concatenate([sc["prior_A"], A]); sc["prior_b"] = np.concatenate([sc["prior_b"], b])
    return sc


def gpu(sc, **opts):
    s = SBASolver(**opts)
    arrays = synthetic.load_into(s, sc)
    return s, arrays


def check_reduced_system(s, p, radius=1e4):
    """The explicit tiles (incl. the border of the pseudo-cameras) and the reduced rhs of the first LM iteration against the
    oracle's dense reduced system at the same point, over the live columns."""
    lay = p.layout()
    S, rhs = s.explicit_dense()
    So, ro = p.reduced(radius=radius)
    live = lay["live"]
    assert S.shape == So.shape
    idx = np.flatnonzero(live)
    scale = np.sqrt(np.abs(np.diag(So))[idx])
    A = S[np.ix_(idx, idx)] / np.outer(scale, scale); B = So[np.ix_(idx, idx)] / np.outer(scale, scale)
    assert np.abs(A - B).max() < 1e-9, np.abs(A - B).max()
    assert relerr(rhs[idx] / scale, ro[idx] / scale) < 1e-9


def run_case(oracle, sc, iters=5, check_S=True, tol_pcg=4, **opts):
    p = oracle.ExtProblem(sc)
    s, (raxis, t, pts) = gpu(sc, **opts)
    s.finalize()
    lay = p.layout()
    info = s.shared_info()
    assert info["n_ext"] == lay["n_ext"] and list(info["model_block"]) == list(lay["model_block"])
    if check_S:
        s1, _ = gpu(sc, **opts)
        s1.solve(max_iterations=1, **FIXED_ITERS)
        check_reduced_system(s1, oracle.ExtProblem(sc))
        s1.close()
    so = p.solve(max_iterations=iters, num_threads=1, **FIXED_ITERS)
    out = s.solve(max_iterations=iters, **FIXED_ITERS)
    assert out["iterations"] == so["iterations"] == iters
    for a, b in zip(out["cost_history"], so["cost_history"]):
        assert abs(a - b) <= TOL_COST * max(b, 1.0), (out["cost_history"], so["cost_history"])
    # same CG on the same system; another summation order moves the exit at 1e-10 (a plateau of the residual) by a few iterations:
    # gate max(4, 8 %) -- a wrong operator or preconditioner changes the counts by far more
    assert all(abs(a - b) <= max(tol_pcg, 0.08 * b) for a, b in zip(out["pcg_iters_history"][1:], so["pcg_iters_history"][1:])), (out["pcg_iters_history"], so["pcg_iters_history"])
    assert relerr(raxis, p.cam_r.reshape(-1, 3)) < 1e-7 and relerr(t, p.cam_t.reshape(-1, 3)) < 1e-7
    return s, p, out, so


@pytest.mark.parametrize("shape,kw", [("small", dict()), ("small", dict(mixed=True)), ("ladybug", dict())])
@pytest.mark.parametrize("mask", [0x10, 0x00, 0x1c, 0x13])
def test_free_intrinsics_lm(oracle, shape, kw, mask):
    """mask: 0x10 = f, pp, k, t free (B constant); 0x00 = everything free (B's step is zero anyway); 0x1c = only f, pp free;
    0x13 = only k, t free."""
    rng = np.random.default_rng(3)
    sc = synthetic.make_scene(shape, **kw)
    sc["intrinsics"] = perturbed_intrinsics(sc, rng)
    sc["intr_fixed_mask"] = np.array([mask], np.uint8)
    s, p, out, so = run_case(oracle, sc)
    assert out["cost_history"][-1] < out["cost_history"][0]
    for c in range(8):
        assert abs(s.intrinsics[0, c] - p.intr.reshape(-1, 10)[0, c]) <= 1e-6 * max(abs(p.intr.reshape(-1, 10)[0, c]), 1e-3), c
    changed = np.abs(s.intrinsics[0] - sc["intrinsics"][0]) > 0
    blk = [0, 1, 1, 2, 2, 2, 3, 3, 4, 4]
    for c in range(10):
        assert changed[c] == (not (mask >> blk[c] & 1) and c < 8), c      # constant blocks (and B) did not move


def test_free_intrinsics_all_constant_equals_base_path(oracle):
    sc = synthetic.make_scene("small")
    base, _ = gpu(sc)
    a = base.solve(max_iterations=4, **FIXED_ITERS)
    sc2 = dict(sc); sc2["intr_fixed_mask"] = np.array([0x1f], np.uint8)
    s, _ = gpu(sc2)
    b = s.solve(max_iterations=4, **FIXED_ITERS)
    assert s.shared_info()["n_groups"] == 0
    assert a["cost_history"] == b["cost_history"]


def test_free_intrinsics_priors_and_two_models(oracle):
    """Two camera models (cameras alternate), the first with every block free and FixedSizeNormalPrior<1,1> / <2,2> / <3,3>
    priors (full matrices), the second with only f free; a heavy point (> 32 views) so the windowed W path runs."""
    rng = np.random.default_rng(8)
    sc = synthetic.make_scene(None, n_cams=40, n_pts=1500, n_obs=7500, seed=4)
    R = synthetic.axis_angle_to_rotmat(sc["raxis_true"])
    cams, uvs = [], []
    have = set(sc["obs_cam"][sc["obs_pt"] == 3].tolist())
    for c in range(sc["n_cams"]):
        if c in have:
            continue
        Pb = R[c].T @ (sc["pts_true"][3] - sc["t_true"][c])
        cams.append(c); uvs.append(synthetic.uv_from_normalized((Pb[:2] / Pb[2])[None], sc["intrinsics"][0])[0])
    sc["obs_cam"] = np.ascontiguousarray(np.concatenate([sc["obs_cam"], np.array(cams, np.int32)]))
    sc["obs_pt"] = np.ascontiguousarray(np.concatenate([sc["obs_pt"], np.full(len(cams), 3, np.int32)]))
    sc["obs_uv"] = np.ascontiguousarray(np.concatenate([sc["obs_uv"], np.array(uvs)])); sc["n_obs"] = sc["obs_cam"].size
    assert np.bincount(sc["obs_pt"]).max() == 40
    truth = sc["intrinsics"][0].copy()
    sc["intrinsics"] = perturbed_intrinsics(sc, rng, 2)
    sc["cam_model"] = (np.arange(sc["n_cams"]) % 2).astype(np.int32)
    sc["intr_fixed_mask"] = np.array([0x00, 0x1e], np.uint8)
    A = np.zeros((4, 9)); b = np.zeros((4, 3))
    A[0, 0] = 1 / 5.0; b[0, 0] = truth[0]                                        # f
    A[1, :4] = (np.eye(2) / 2.0 + rng.normal(0, 0.05, (2, 2))).ravel(); b[1, :2] = truth[1:3]      # pp
    A[2] = (np.eye(3) * 50.0 + rng.normal(0, 3.0, (3, 3))).ravel(); b[2] = truth[3:6]           # k
    A[3, 0] = 1 / 3.0; b[3, 0] = truth[0]                                        # f of the second model
    sc["intr_prior_model"] = np.array([0, 0, 0, 1], np.int32); sc["intr_prior_block"] = np.array([0, 1, 2, 0], np.int32)
    sc["intr_prior_A"] = A; sc["intr_prior_b"] = b
    s, p, out, so = run_case(oracle, sc, iters=6)
    assert relerr(s.intrinsics, p.intr.reshape(-1, 10)) < 1e-7
    assert np.array_equal(s.intrinsics[1, 1:], sc["intrinsics"][1, 1:]) and s.intrinsics[1, 0] != sc["intrinsics"][1, 0]


@pytest.mark.parametrize("kw", [dict(), dict(mixed=True)])
@pytest.mark.parametrize("operator_mode", [3, 2])
def test_constant_sensor_offsets_lm(oracle, kw, operator_mode):
    """PoseTransformCost inside the LM (explicit tiles and the implicit operator with the factored camera pass)."""
    sc = add_rig(synthetic.make_scene("small", **kw), np.random.default_rng(5), offsets=True)
    p = oracle.ExtProblem(sc)
    eo = p.evaluate()
    s, _ = gpu(sc, operator_mode=operator_mode)
    e = s.evaluate()
    assert e["ok"] and eo["ok"]
    assert relerr(e["residuals"], eo["residuals"]) < 1e-12
    assert relerr(e["jac_cam"], eo["jac_cam"]) < 1e-12 and relerr(e["jac_pt"], eo["jac_pt"]) < 1e-12
    s.close()
    run_case(oracle, sc, operator_mode=operator_mode, check_S=operator_mode == 3)


@pytest.mark.parametrize("offsets", [False, True])
def test_lever_arms_lm(oracle, offsets):
    """LeverArmCost / PoseTransformLeverArmCost: two mountings (one free, one constant), cameras without a mounting in between."""
    sc = add_rig(synthetic.make_scene("small", mixed=True), np.random.default_rng(6), offsets=offsets, n_mounts=2)
    sc["mount_fixed"] = np.array([0, 1], np.uint8)
    sc["mount_b"][1], sc["mount_l"][1] = sc["mount_b_true"][1], sc["mount_l_true"][1]
    sc = with_pose_priors(sc)
    s, p, out, so = run_case(oracle, sc, iters=6)
    assert relerr(s.mount_b, p.mount_b.reshape(-1, 3)) < 1e-7 and relerr(s.mount_l, p.mount_l.reshape(-1, 3)) < 1e-7
    assert np.array_equal(s.mount_b[1], sc["mount_b"][1])
    assert np.abs(s.mount_l[0] - sc["mount_l_true"][0]).max() < np.abs(sc["mount_l"][0] - sc["mount_l_true"][0]).max()


def test_everything_free_ladybug(oracle):
    """Free intrinsics + two free mountings + offsets + stiffness matrices on the Ladybug shape (the cross term of a model and a
    mounting on the same observation)."""
    rng = np.random.default_rng(11)
    sc = add_rig(synthetic.make_scene("ladybug"), rng, offsets=True, n_mounts=2)
    sc["intrinsics"] = perturbed_intrinsics(sc, rng)
    sc["intr_fixed_mask"] = np.array([0x10], np.uint8)
    sc["obs_info"] = rng.normal(0, 0.1, (sc["n_obs"], 2, 2)) + np.eye(2)
    sc = with_pose_priors(sc)
    s, p, out, so = run_case(oracle, sc, iters=5, tol_pcg=15)   # ~170 CG iterations on a plateau: the exit moves with the summation order
    assert relerr(s.intrinsics, p.intr.reshape(-1, 10)) < 1e-7 and relerr(s.mount_l, p.mount_l.reshape(-1, 3)) < 1e-7


def test_shared_blocks_need_the_explicit_system():
    sc = synthetic.make_scene("small")
    sc["intr_fixed_mask"] = np.array([0x10], np.uint8)
    s, _ = gpu(sc, operator_mode=2)
    with pytest.raises(SBAError) as err:
        s.finalize()
    assert err.value.code == -3


def test_free_intrinsics_deterministic_rerun():
    sc = synthetic.make_scene("small"); sc["intr_fixed_mask"] = np.array([0x10], np.uint8)
    sc["intrinsics"] = perturbed_intrinsics(sc, np.random.default_rng(1))
    h = []
    for _ in range(2):
        s, _ = gpu(sc)
        h.append(s.solve(max_iterations=4, **FIXED_ITERS)["cost_history"])
        s.close()
    assert h[0] == h[1]
