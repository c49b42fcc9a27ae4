"""BAL loader (sba_bal_* in include/sba_b200.h): file round trip, conversion to the engine's conventions checked against
a numpy statement of the BAL camera model (modules/g2o/g2o/examples/bal/bal_example.cpp:168-204) and against the
oracle's cost functors, error behaviour.  CPU tests need only the shared library; the GPU test runs the loaded problem."""
import os
import sys

import numpy as np
import pytest

from steviapp_b200 import SBASolver, bal, synthetic

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))

F = np.diag([1.0, -1.0, -1.0])


def rodrigues(r):
    th = np.linalg.norm(r)
    K = np.array([[0, -r[2], r[1]], [r[2], 0, -r[0]], [-r[1], r[0], 0]])
    if th < 1e-8:
        return np.eye(3) + K
    return np.eye(3) + np.sin(th) / th * K + (1 - np.cos(th)) / th ** 2 * K @ K


def bal_project(cam9, X):
    """bal_example.cpp:168-204: P = R X + t, p = -P.xy / P.z, pixel = f (1 + k1 |p|^2 + k2 |p|^4) p."""
    P = rodrigues(cam9[:3]) @ X + cam9[3:6]
    p = -P[:2] / P[2]
    r2 = p @ p
    return cam9[6] * (1 + cam9[7] * r2 + cam9[8] * r2 * r2) * p, P


def make_bal(seed=7, n_cams=6, n_pts=60, noise=0.0):
    rng = np.random.default_rng(seed)
    cam9 = np.zeros((n_cams, 9))
    cam9[:, :3] = rng.normal(0, 0.25, (n_cams, 3))
    cam9[0, :3] = 0.0            # Taylor branch of the Rodrigues formula; converted pose is exactly a pi rotation
    cam9[1, :3] = [0.0, 0.0, 3.0]   # large roll
    cam9[:, 3:6] = rng.normal(0, 0.5, (n_cams, 3))
    cam9[:, 6] = rng.uniform(600, 1500, n_cams)
    cam9[:, 7] = rng.uniform(-0.05, 0.02, n_cams)
    cam9[:, 8] = rng.uniform(-0.005, 0.01, n_cams)
    xyz = np.column_stack([rng.uniform(-3, 3, n_pts), rng.uniform(-3, 3, n_pts), rng.uniform(-14, -8, n_pts)])
    ci, pi, xy = [], [], []
    for p in range(n_pts):
        for c in rng.choice(n_cams, size=rng.integers(2, n_cams + 1), replace=False):
            px, P = bal_project(cam9[c], xyz[p])
            if P[2] < -1.0:
                ci.append(c); pi.append(p); xy.append(px + rng.normal(0, noise, 2) if noise else px)
    return np.array(ci, np.int32), np.array(pi, np.int32), np.array(xy), cam9, xyz


def test_write_read_round_trip_is_bit_exact(tmp_path):
    ci, pi, xy, cam9, xyz = make_bal()
    path = tmp_path / "problem.txt"
    bal.write_bal(path, ci, pi, xy, cam9, xyz)
    raw = bal.read_bal_raw(path)
    assert (raw["n_cams"], raw["n_pts"], raw["n_obs"]) == (cam9.shape[0], xyz.shape[0], ci.shape[0])
    assert np.array_equal(raw["cam_idx"], ci) and np.array_equal(raw["pt_idx"], pi)
    assert np.array_equal(raw["xy"], xy) and np.array_equal(raw["cam9"], cam9) and np.array_equal(raw["xyz"], xyz)


def test_reads_the_layout_the_reference_example_reads(tmp_path):
    """Free-form whitespace, exponents, one value per line for cameras and points (the published BAL files)."""
    path = tmp_path / "tiny.txt"
    path.write_text("2 1   2\n0 0   -3.5e+00 1.25\n1 0 2 -1\n" + "\n".join(str(v) for v in range(18)) + "\n1.0\n2.0\n-3e0\n")
    raw = bal.read_bal_raw(path)
    assert np.array_equal(raw["xy"], [[-3.5, 1.25], [2.0, -1.0]])
    assert np.array_equal(raw["cam9"], np.arange(18.0).reshape(2, 9))
    assert np.array_equal(raw["xyz"], [[1.0, 2.0, -3.0]])


def test_conversion_matches_the_bal_camera_model(tmp_path):
    ci, pi, xy, cam9, xyz = make_bal()
    path = tmp_path / "problem.txt"
    bal.write_bal(path, ci, pi, xy, cam9, xyz)
    sc = bal.load_bal(path)
    assert sc["undistort_failed"] == 0
    assert np.array_equal(sc["cam_model"], np.arange(cam9.shape[0])) and np.array_equal(sc["intrinsics"][:, 0], cam9[:, 6])
    assert not sc["intrinsics"][:, 1:].any() and np.array_equal(sc["pts"], xyz)
    assert np.all(np.linalg.norm(sc["raxis"], axis=1) <= np.pi + 1e-12)
    worst_P = worst_res = worst_redist = 0.0
    for k in range(ci.shape[0]):
        c = ci[k]
        Re = rodrigues(sc["raxis"][c])
        Pbar = Re.T @ (xyz[pi[k]] - sc["t"][c])            # the engine's sensor-frame point (sensor -> world pose)
        _, P = bal_project(cam9[c], xyz[pi[k]])
        worst_P = max(worst_P, np.abs(Pbar - F @ P).max())
        assert Pbar[2] > 0                                  # in front of a camera that looks down +z
        f = cam9[c, 6]
        worst_res = max(worst_res, np.abs(f * Pbar[:2] / Pbar[2] - sc["obs_uv"][k]).max())   # plane cost, k = 0
        # distorting the converted pixel again with the file's model gives the measurement back
        pu = np.array([sc["obs_uv"][k, 0], -sc["obs_uv"][k, 1]]) / f
        r2 = pu @ pu
        worst_redist = max(worst_redist, np.abs(f * (1 + cam9[c, 7] * r2 + cam9[c, 8] * r2 * r2) * pu - xy[k]).max())
    assert worst_P < 1e-12 and worst_res < 1e-9 and worst_redist < 1e-10, (worst_P, worst_res, worst_redist)


def test_oracle_cost_is_zero_on_an_exact_bal_problem(tmp_path):
    """The converted problem through the oracle's restatement of the reference cost functors (not the numpy formula above)."""
    import oracle_py
    ci, pi, xy, cam9, xyz = make_bal(seed=11)
    path = tmp_path / "problem.txt"
    bal.write_bal(path, ci, pi, xy, cam9, xyz)
    sc = bal.load_bal(path)
    ev = oracle_py.Problem(sc).evaluate()
    assert ev["cost"] < 1e-16 * ci.shape[0] * 1e6, ev["cost"]


def test_bal_cost_in_the_files_own_metric(tmp_path):
    """sba_bal_cost = 1/2 sum |bal_project - xy|^2 (the numpy statement above), for the file's parameters and -- through
    sba_bal_cameras_from_engine -- for engine-convention poses; the round trip file -> engine -> file reproduces the cameras."""
    ci, pi, xy, cam9, xyz = make_bal(seed=5, noise=0.7)
    path = tmp_path / "problem.txt"
    bal.write_bal(path, ci, pi, xy, cam9, xyz)
    want = 0.5 * sum(((bal_project(cam9[c], xyz[p])[0] - m) ** 2).sum() for c, p, m in zip(ci, pi, xy))
    cost, rms = bal.bal_cost(path)
    assert abs(cost - want) <= 1e-12 * want and abs(rms - np.sqrt(2 * want / (2 * ci.size))) <= 1e-12 * rms
    sc = bal.load_bal(path)
    cost2, _ = bal.bal_cost(path, sc["raxis"], sc["t"], sc["pts"])          # engine conventions back to BAL cameras
    assert abs(cost2 - want) <= 1e-10 * want
    moved = sc["pts"] + 0.01
    cost3, _ = bal.bal_cost(path, sc["raxis"], sc["t"], moved)
    want3 = 0.5 * sum(((bal_project(cam9[c], xyz[p] + 0.01)[0] - m) ** 2).sum() for c, p, m in zip(ci, pi, xy))
    assert abs(cost3 - want3) <= 1e-10 * want3


def test_undistortion_failures_are_surfaced(tmp_path):
    """A wide-angle pixel beyond the turning point of 1 + k1 r^2 + k2 r^4 has no undistorted radius: counted AND warned about."""
    ci, pi, xy, cam9, xyz = make_bal(seed=3)
    cam9[:, 7] = -0.6; cam9[:, 8] = 0.0
    xy = xy.copy(); xy[0] = cam9[ci[0], 6] * np.array([2.0, 2.0])           # normalised radius 2.8: past the maximum of r (1 - 0.6 r^2)
    path = tmp_path / "problem.txt"
    bal.write_bal(path, ci, pi, xy, cam9, xyz)
    with pytest.warns(RuntimeWarning, match="could not be removed"):
        sc = bal.load_bal(path)
    assert sc["undistort_failed"] >= 1


def test_loaded_problem_shards_by_points(tmp_path):
    """SURVEY 8(e) sharding applies to a loaded file as to the synthetic scenes: contiguous point ranges, every observation
    on exactly one rank, cameras replicated."""
    ci, pi, xy, cam9, xyz = make_bal(seed=2, n_cams=8, n_pts=300)
    path = tmp_path / "problem.txt"
    bal.write_bal(path, ci, pi, xy, cam9, xyz)
    sc = bal.load_bal(path)
    parts = [synthetic.shard_scene(sc, r, 3) for r in range(3)]
    assert sum(q["n_pts"] for q in parts) == sc["n_pts"] and sum(q["n_obs"] for q in parts) == sc["n_obs"]
    for q in parts:
        lo, hi = q["point_range"]
        assert np.array_equal(q["pts"], sc["pts"][lo:hi]) and np.array_equal(q["raxis"], sc["raxis"])
        assert q["obs_pt"].min() >= 0 and q["obs_pt"].max() < q["n_pts"]
    obs = max(q["n_obs"] for q in parts)
    assert obs <= 1.2 * sc["n_obs"] / 3 + 10          # balanced by observation count


@pytest.mark.parametrize("text, what", [
    ("", "bad header"), ("0 5 5\n", "bad header"), ("2 2\n", "bad header"),
    ("1 1 1\n0 3 1.0 2.0\n" + "0 " * 12, "out of bounds"), ("1 1 1\n-1 0 1.0 2.0\n" + "0 " * 12, "out of bounds"),
    ("1 1 1\n0 0 1.0 2.0\n" + "0 " * 11, "ends early"), ("1 1 1\n0 0 1.0 x\n" + "0 " * 12, "not a number"),
])
def test_malformed_files_are_rejected(tmp_path, text, what):
    path = tmp_path / "bad.txt"
    path.write_text(text)
    with pytest.raises(ValueError, match=what):
        bal.read_bal_raw(path)


def test_missing_file():
    with pytest.raises(ValueError, match="cannot read"):
        bal.load_bal("/nonexistent/problem.txt")


@pytest.mark.gpu
def test_bal_problem_on_the_gpu(tmp_path):
    """Exact file: zero cost on the device.  Noisy file with perturbed points: the device residuals equal the oracle's on the
    converted problem to 1e-12, and the LM brings the cost down."""
    import oracle_py
    ci, pi, xy, cam9, xyz = make_bal(seed=3, n_cams=8, n_pts=200)
    path = tmp_path / "exact.txt"
    bal.write_bal(path, ci, pi, xy, cam9, xyz)
    with SBASolver() as s:
        synthetic.load_into(s, bal.load_bal(path))
        assert s.evaluate(jacobians=False)["cost"] < 1e-12
    rng = np.random.default_rng(5)
    path = tmp_path / "noisy.txt"
    bal.write_bal(path, ci, pi, xy + rng.normal(0, 0.5, xy.shape), cam9, xyz + rng.normal(0, 0.05, xyz.shape))
    sc = bal.load_bal(path)
    ref = oracle_py.Problem(sc).evaluate()
    with SBASolver(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0) as s:
        synthetic.load_into(s, sc)
        ev = s.evaluate()
        assert abs(ev["cost"] - ref["cost"]) <= 1e-12 * ref["cost"]
        assert np.abs(ev["residuals"] - ref["residuals"]).max() <= 1e-12 * max(np.abs(ref["residuals"]).max(), 1.0)
        r = s.solve(max_iterations=8)
        assert r["final_cost"] < 0.2 * r["initial_cost"]
