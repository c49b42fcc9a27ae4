"""SURVEY 8(f) row 2 on the GPU: UV <-> UV cross-projection factors (UV2UVCost, costfunctors/modularuvprojection.h:660-795)
with Huber / Cauchy / Arctan losses (sbamodules/correspondencessetsbamodule.cpp:2459-2471) against the extended oracle."""
import numpy as np
import pytest

from conftest import relerr
from steviapp_b200 import SBASolver, synthetic
from test_pairs_host import cross_projector_scene, image_pair_scene, pair_scene

pytestmark = pytest.mark.gpu

TOL_COST = 1e-8
FIXED_ITERS = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)


def gpu(sc, **opts):
    s = SBASolver(**opts)
    arrays = synthetic.load_into(s, sc)
    return s, arrays


def test_reference_known_answer_image_pair():
    # tests/test_sba_solver/main.cpp:335-492: the factor graph of two images and five correspondences costs < 1e-6 at the truth
    s, _ = gpu(image_pair_scene())
    e = s.evaluate(residuals=False, jacobians=False)
    r, j = s.evaluate_pairs()
    assert abs(e["cost"]) < 1e-6 and np.abs(r).max() < 1e-9


def test_reference_known_answer_cross_projector(sba_lib):
    # tests/test_uv_projectors/main.cpp:258-382: frame camera x push-broom camera, cost < 1e-6 at the truth
    s, _ = gpu(cross_projector_scene(sba_lib))
    e = s.evaluate(residuals=False, jacobians=False)
    r, j = s.evaluate_pairs()
    assert abs(e["cost"]) < 1e-6 and np.abs(r).max() < 1e-9


@pytest.mark.parametrize("with_info", [False, True])
def test_pair_residuals_and_jacobians(oracle, with_info):
    sc = pair_scene(n_pairs=500)
    if with_info:
        rng = np.random.default_rng(2)
        sc["pair_info"] = (rng.normal(0, 0.2, (500, 2, 2, 2)) + np.eye(2)).reshape(500, 8)
    eo = oracle.ExtProblem(sc).evaluate()
    s, _ = gpu(sc)
    r, j = s.evaluate_pairs()
    assert relerr(r, eo["pair_residuals"]) < 1e-12
    assert relerr(j, eo["pair_jac"][:, :, :12]) < 1e-12
    e = s.evaluate(residuals=False, jacobians=False)
    assert abs(e["cost"] - eo["cost"]) <= 1e-12 * eo["cost"]


@pytest.mark.parametrize("loss", [None, "huber", "cauchy", "arctan"])
def test_pairs_lm_with_losses(oracle, loss):
    """Projection observations + 600 factors, 10 % of them gross outliers; the LM cost history against the oracle."""
    sc = pair_scene(n_pairs=600, outliers=0.1)
    sc["loss"] = loss
    p = oracle.ExtProblem(sc)
    so = p.solve(max_iterations=5, num_threads=1, **FIXED_ITERS)
    s, (raxis, t, pts) = gpu(sc)
    out = s.solve(max_iterations=5, **FIXED_ITERS)
    for a, b in zip(out["cost_history"], so["cost_history"]):
        assert abs(a - b) <= TOL_COST * max(b, 1.0), (out["cost_history"], so["cost_history"])
    assert all(abs(a - b) <= max(4, 0.08 * b) for a, b in zip(out["pcg_iters_history"][1:], so["pcg_iters_history"][1:]))
    assert relerr(raxis, p.cam_r.reshape(-1, 3)) < 1e-7 and relerr(t, p.cam_t.reshape(-1, 3)) < 1e-7


def test_pairs_only_relative_pose(oracle):
    """No landmark at all: two fixed cameras and six free ones tied by factors only (a camera pair without a common point gets
    its block of the reduced system from the factors alone)."""
    sc = pair_scene(n_pairs=700, noise=0.5)
    keep = np.zeros(sc["n_obs"], bool)                 # drop every projection observation
    for k in ("obs_cam", "obs_pt", "obs_uv"):
        sc[k] = np.ascontiguousarray(sc[k][keep])
    sc["n_obs"] = 0
    sc["prior_kind"] = np.zeros(0, np.int32); sc["prior_index"] = np.zeros(0, np.int32); sc["prior_A"] = np.zeros((0, 9)); sc["prior_b"] = np.zeros((0, 3))
    sc["pt_fixed"] = np.ones(sc["n_pts"], np.uint8)
    sc["cam_fixed"] = sc["cam_fixed"].copy(); sc["cam_fixed"][1] = 1; sc["raxis"][1] = sc["raxis_true"][1]; sc["t"][1] = sc["t_true"][1]
    p = oracle.ExtProblem(sc)
    so = p.solve(max_iterations=6, num_threads=1, **FIXED_ITERS)
    s, (raxis, t, pts) = gpu(sc)
    out = s.solve(max_iterations=6, **FIXED_ITERS)
    for a, b in zip(out["cost_history"], so["cost_history"]):
        assert abs(a - b) <= TOL_COST * max(b, 1.0), (out["cost_history"], so["cost_history"])
    assert out["cost_history"][-1] < 0.05 * out["cost_history"][0]


def test_pairs_with_free_intrinsics(oracle):
    """The factors' Jacobians with respect to the (shared, free) intrinsic blocks enter the border of the reduced system."""
    sc = pair_scene(n_pairs=500)
    rng = np.random.default_rng(4)
    intr = sc["intrinsics"].copy(); intr[0, 0] *= 1.003; intr[0, 1:3] += rng.normal(0, 2.0, 2); intr[0, 3] += 0.003
    sc["intrinsics"] = intr
    sc["intr_fixed_mask"] = np.array([0x10], np.uint8)
    sc["loss"] = "huber"
    p = oracle.ExtProblem(sc)
    so = p.solve(max_iterations=5, num_threads=1, **FIXED_ITERS)
    s, _ = gpu(sc)
    out = s.solve(max_iterations=5, **FIXED_ITERS)
    for a, b in zip(out["cost_history"], so["cost_history"]):
        assert abs(a - b) <= TOL_COST * max(b, 1.0), (out["cost_history"], so["cost_history"])
    assert relerr(s.intrinsics, p.intr.reshape(-1, 10)) < 1e-7
