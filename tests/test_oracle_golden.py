"""Pins the oracle against every known-answer vector the reference's own tests hold for this path
(SURVEY 8c): LibStevi rotation matrices, the InvertPose convention, the zero-residual vectors of
test_uv_projectors and test_sba_solver::simplePnP.  CPU only."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from conftest import ptr

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def rot(oracle, r, x):
    lib = oracle.load()
    r = np.asarray(r, dtype=np.float64); x = np.asarray(x, dtype=np.float64); out = np.empty(3)
    lib.oracle_angle_axis_rotate(ptr(r), ptr(x), ptr(out))
    return out


# tests/test_geometry/testRotations.cpp:25-65 (rodriguezFormula known matrices, tolerance 1e-5 there)
RODRIGUES = [
    ((0, 0, 0), np.eye(3)),
    ((np.pi / 2, 0, 0), [[1, 0, 0], [0, 0, -1], [0, 1, 0]]),
    ((0, np.pi / 2, 0), [[0, 0, 1], [0, 1, 0], [-1, 0, 0]]),
    ((0, 0, np.pi / 2), [[0, -1, 0], [1, 0, 0], [0, 0, 1]]),
    ((np.pi, 0, 0), [[1, 0, 0], [0, -1, 0], [0, 0, -1]]),
    ((0, np.pi, 0), [[-1, 0, 0], [0, 1, 0], [0, 0, -1]]),
    ((0, 0, np.pi), [[-1, 0, 0], [0, -1, 0], [0, 0, 1]]),
]


@pytest.mark.parametrize("r,M", RODRIGUES)
def test_rodrigues_known_matrices(oracle, r, M):
    R = np.stack([rot(oracle, r, e) for e in np.eye(3)], axis=1)
    assert np.linalg.norm(R.T @ np.asarray(M, dtype=float) - np.eye(3)) < 1e-12


def test_invert_pose_is_inverse(oracle):
    # tests/test_functional_utils/main.cpp:198-218: pose o InvertPose(pose) = identity
    lib = oracle.load()
    rng = np.random.default_rng(0)
    for _ in range(50):
        r = rng.normal(size=3); t = rng.normal(size=3) * 10; x = rng.normal(size=3) * 5
        ri = np.empty(3); ti = np.empty(3)
        lib.oracle_rigid_inverse(ptr(r), ptr(t), ptr(ri), ptr(ti))
        y = rot(oracle, r, x) + t
        back = rot(oracle, ri, y) + ti
        assert np.abs(back - x).max() < 1e-12


def _scene(points, uv, f=35.0, pp=(320.0, 240.0), spherical=False, r=(0, 0, 0), t=(0, 0, 0)):
    n = len(points)
    return {
        "raxis": np.array([r], dtype=float), "t": np.array([t], dtype=float), "pts": np.array(points, dtype=float),
        "cam_model": np.zeros(1, np.int32), "cam_fixed": np.zeros(1, np.uint8),
        "cost_kind": np.array([1 if spherical else 0], np.uint8), "pt_fixed": np.zeros(n, np.uint8),
        "intrinsics": np.array([[f, pp[0], pp[1], 0, 0, 0, 0, 0, 0, 0]], dtype=float),
        "obs_cam": np.zeros(n, np.int32), "obs_pt": np.arange(n, dtype=np.int32),
        "obs_uv": np.array(uv, dtype=float), "obs_info": None,
        "prior_kind": None, "prior_index": None, "prior_A": None, "prior_b": None,
    }


def uv_projector_scene(spherical):
    # tests/test_uv_projectors/main.cpp:37-110 (plane) and :112-186 (spherical):
    # f = 35, pp = (320,240), pose 0, X = (42,33,35), uv = (362,273)
    return _scene([[42, 33, 35]], [[362, 273]], spherical=spherical)


def simple_pnp_scene():
    # tests/test_sba_solver/main.cpp:61-236: five hard-coded points, identity pose, uv = f X.xy/X.z + pp
    pts = np.array([[42, 69, 33], [33, 27, 50], [-6, -99, 70], [-49, -28, 12], [-69, -42, 33]], dtype=float)
    uv = 35.0 * pts[:, :2] / pts[:, 2:3] + np.array([320.0, 240.0])
    return _scene(pts, uv)


@pytest.mark.parametrize("spherical", [False, True])
def test_uv_projector_zero_residual(oracle, spherical):
    e = oracle.Problem(uv_projector_scene(spherical)).evaluate()
    assert e["ok"]
    assert abs(e["cost"]) < 1e-6          # the reference's own assertion
    assert np.abs(e["residuals"]).max() < 1e-12


def test_simple_pnp_zero_cost(oracle):
    e = oracle.Problem(simple_pnp_scene()).evaluate()
    assert e["ok"] and abs(e["cost"]) < 1e-6
    assert np.abs(e["residuals"]).max() < 1e-11


def test_cheirality_failure(oracle):
    # plane cost: functor returns false for Pbar.z < 0 (modularuvprojection.h:395-399); spherical has no test
    sc = _scene([[1, 2, -5]], [[300, 200]])
    assert not oracle.Problem(sc).evaluate()["ok"]
    sc = _scene([[1, 2, -5]], [[300, 200]], spherical=True)
    assert oracle.Problem(sc).evaluate()["ok"]


def test_bs_block_is_unused(oracle):
    # reference quirk: B_dist is built from ts, Bs never enters (pinholecameraprojectormodule.h:46-47)
    uv = np.array([1234.5, 987.25])
    a = np.array([2000.0, 2000, 1500, -0.05, 0.01, 0.001, 1e-4, -1e-4, 0.0, 0.0])
    b = a.copy(); b[8:] = [0.3, -0.2]
    assert np.array_equal(oracle.dir_from_uv(uv, a), oracle.dir_from_uv(uv, b))


def test_forward_model_inverts_dir_from_uv(oracle):
    intr = np.array([2000.0, 2000, 1500, -0.05, 0.01, 0.0, 1e-4, -1e-4, 0, 0])
    rng = np.random.default_rng(3)
    for _ in range(100):
        xn = rng.uniform(-0.6, 0.6, 2)
        uv = oracle.uv_from_dir(xn, intr)
        d = oracle.dir_from_uv(uv, intr)
        assert np.abs(d[:2] / d[2] - xn).max() < 1e-13


def test_golden_fixture_matches_oracle(oracle):
    """tests/golden/small_scene.json was written by tests/golden/make_golden.py from this oracle;
    guards the oracle (and the generator) against silent drift."""
    with open(os.path.join(GOLD, "small_scene.json")) as f:
        g = json.load(f)
    from steviapp_b200 import synthetic
    sc = synthetic.make_scene(**g["scene_args"])
    p = oracle.Problem(sc)
    e = p.evaluate()
    assert abs(e["cost"] - g["initial_cost"]) <= 1e-12 * g["initial_cost"]
    idx = g["sample_obs"]
    assert np.abs(e["residuals"][idx] - np.array(g["residuals"])).max() < 1e-9
    assert np.abs(e["jac_cam"][idx] - np.array(g["jac_cam"])).max() < 1e-9
    s = p.solve(max_iterations=g["lm_iterations"], function_tolerance=0.0, parameter_tolerance=0.0,
                gradient_tolerance=0.0, num_threads=1)
    assert abs(s["final_cost"] - g["final_cost"]) <= 1e-9 * g["final_cost"]
