"""Parity of the CUDA path (through the C ABI) against the CPU oracle, on identical seeded inputs.

Gates (north star): residuals and Jacobians 1e-12 relative; final cost after a fixed number of LM
iterations 1e-8 relative.  Everything here needs a B200 (marker gpu)."""
import json
import os

import numpy as np
import pytest

from conftest import relerr
from steviapp_b200 import SBASolver, SBAError, synthetic
from test_oracle_golden import uv_projector_scene, simple_pnp_scene, GOLD

pytestmark = pytest.mark.gpu

TOL_J = 1e-12
TOL_COST = 1e-8
FIXED_ITERS = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)


def gpu_solver(sc, **opts):
    s = SBASolver(**opts)
    arrays = synthetic.load_into(s, _fill(sc))
    return s, arrays


def _fill(sc):
    sc = dict(sc)
    n_c = sc["raxis"].shape[0]
    sc.setdefault("prior_kind", None)
    if sc.get("prior_kind") is None:
        sc["prior_kind"] = np.zeros(0, np.int32); sc["prior_index"] = np.zeros(0, np.int32)
        sc["prior_A"] = np.zeros((0, 9)); sc["prior_b"] = np.zeros((0, 3))
    return sc


@pytest.mark.parametrize("spherical", [False, True])
def test_reference_known_answer_uv_projector(spherical):
    # tests/test_uv_projectors/main.cpp:37-186: cost < 1e-6 at the exact solution
    s, _ = gpu_solver(uv_projector_scene(spherical))
    e = s.evaluate()
    assert e["ok"] and abs(e["cost"]) < 1e-6 and np.abs(e["residuals"]).max() < 1e-12


def test_reference_known_answer_simple_pnp():
    # tests/test_sba_solver/main.cpp:219-234
    s, _ = gpu_solver(simple_pnp_scene())
    e = s.evaluate()
    assert e["ok"] and abs(e["cost"]) < 1e-6


@pytest.mark.parametrize("kw", [dict(), dict(spherical=True), dict(mixed=True)])
@pytest.mark.parametrize("with_info", [False, True])
def test_residuals_jacobians(oracle, kw, with_info):
    sc = synthetic.make_scene("small", **kw)
    if with_info:
        sc["obs_info"] = np.random.default_rng(5).normal(size=(sc["n_obs"], 2, 2)) + 2 * np.eye(2)
    o = oracle.Problem(sc).evaluate()
    s, _ = gpu_solver(sc)
    e = s.evaluate()
    assert e["ok"]
    assert relerr(e["residuals"], o["residuals"]) < TOL_J
    assert relerr(e["jac_cam"], o["jac_cam"]) < TOL_J
    assert relerr(e["jac_pt"], o["jac_pt"]) < TOL_J
    assert abs(e["cost"] - o["cost"]) <= 1e-12 * o["cost"]


def test_residuals_jacobians_ladybug(oracle):
    sc = synthetic.make_scene("ladybug")
    o = oracle.Problem(sc).evaluate()
    s, _ = gpu_solver(sc)
    e = s.evaluate()
    assert relerr(e["residuals"], o["residuals"]) < TOL_J
    assert relerr(e["jac_cam"], o["jac_cam"]) < TOL_J
    assert relerr(e["jac_pt"], o["jac_pt"]) < TOL_J


def test_golden_fixture():
    with open(os.path.join(GOLD, "small_scene.json")) as f:
        g = json.load(f)
    sc = synthetic.make_scene(**g["scene_args"])
    s, _ = gpu_solver(sc)
    e = s.evaluate()
    idx = g["sample_obs"]
    assert abs(e["cost"] - g["initial_cost"]) <= 1e-12 * g["initial_cost"]
    assert relerr(e["residuals"][idx], np.array(g["residuals"])) < TOL_J
    assert relerr(e["jac_cam"][idx], np.array(g["jac_cam"])) < TOL_J
    assert relerr(e["jac_pt"][idx], np.array(g["jac_pt"])) < TOL_J


def test_cheirality_reported(oracle):
    sc = synthetic.make_scene("tiny")
    sc["pts"] = sc["pts"].copy()
    sc["pts"][0] = sc["t"][sc["obs_cam"][sc["obs_pt"] == 0][0]] * 3.0
    o = oracle.Problem(sc).evaluate()
    s, _ = gpu_solver(sc)
    e = s.evaluate()
    assert not o["ok"] and not e["ok"]
    assert relerr(e["residuals"], o["residuals"]) < TOL_J     # failed rows are zero on both sides
    with pytest.raises(SBAError) as err:
        s.solve(max_iterations=2)
    assert err.value.code == -4


@pytest.mark.parametrize("kw", [dict(), dict(mixed=True)])
def test_normal_blocks(oracle, kw):
    sc = synthetic.make_scene("small", **kw)
    sc["pt_fixed"] = sc["pt_fixed"].copy(); sc["pt_fixed"][5] = 1
    o = oracle.Problem(sc).normal_blocks()
    s, _ = gpu_solver(sc)
    g = s.normal_blocks()
    for k in ("gc", "gp", "Hcc", "Hpp"):
        assert relerr(g[k], o[k]) < 1e-11, k


@pytest.mark.parametrize("kw", [dict(), dict(spherical=True), dict(mixed=True)])
@pytest.mark.parametrize("operator_mode", [0, 1, 2, 3])
def test_schur_matvec_and_rhs(oracle, kw, operator_mode):
    """operator_mode 0 = auto, 1 = implicit with materialised A, C, 2 = implicit with the factored camera pass
    (A, C rebuilt from Pbar), 3 = explicit reduced camera system (dense tiles, static contribution slots)."""
    sc = synthetic.make_scene("small", **kw)
    rng = np.random.default_rng(9)
    lam_c = rng.uniform(0.5, 2.0, (sc["n_cams"], 6)); lam_p = rng.uniform(0.5, 2.0, (sc["n_pts"], 3))
    x = rng.normal(size=(sc["n_cams"], 6))
    yo, ro = oracle.Problem(sc).schur_matvec(lam_c, lam_p, x)
    s, _ = gpu_solver(sc, operator_mode=operator_mode)
    y, r = s.schur_matvec(lam_c, lam_p, x)
    assert relerr(y, yo) < 1e-11
    assert relerr(r, ro) < 1e-11


def test_operator_modes_agree_on_lm():
    sc = synthetic.make_scene("ladybug", mixed=True)
    b = gpu_solver(sc, operator_mode=1)[0].solve(max_iterations=4, **FIXED_ITERS)
    for mode in (0, 2, 3):
        a = gpu_solver(sc, operator_mode=mode)[0].solve(max_iterations=4, **FIXED_ITERS)
        assert abs(a["final_cost"] - b["final_cost"]) <= 1e-10 * b["final_cost"], mode
        for ca, cb in zip(a["cost_history"], b["cost_history"]):
            assert abs(ca - cb) <= 1e-9 * cb, mode


def test_explicit_system_with_fixed_cameras_and_heavy_points(oracle):
    """Explicit tiles: a fixed camera in the middle, a point seen by more than 32 cameras, more than one tile per side."""
    sc = synthetic.make_scene(None, n_cams=150, n_pts=600, n_obs=6000, seed=5, window=60)
    sc["cam_fixed"] = sc["cam_fixed"].copy(); sc["cam_fixed"][77] = 1
    rng = np.random.default_rng(3)
    lam_c = rng.uniform(0.5, 2.0, (sc["n_cams"], 6)); lam_p = rng.uniform(0.5, 2.0, (sc["n_pts"], 3))
    x = rng.normal(size=(sc["n_cams"], 6))
    yo, ro = oracle.Problem(sc).schur_matvec(lam_c, lam_p, x)
    for mode in (2, 3):
        s, _ = gpu_solver(sc, operator_mode=mode)
        y, r = s.schur_matvec(lam_c, lam_p, x)
        assert relerr(y, yo) < 1e-11, mode
        assert relerr(r, ro) < 1e-11, mode


@pytest.mark.parametrize("name", ["small_scene", "small_scene_spherical", "small_scene_mixed", "ladybug_scene"])
def test_lm_final_cost_golden(name):
    """Final cost after a fixed number of LM iterations against the committed oracle fixture."""
    with open(os.path.join(GOLD, name + ".json")) as f:
        g = json.load(f)
    sc = synthetic.make_scene(**g["scene_args"])
    s, _ = gpu_solver(sc)
    out = s.solve(max_iterations=g["lm_iterations"], **FIXED_ITERS)
    assert out["iterations"] == g["lm_iterations"]
    assert abs(out["final_cost"] - g["final_cost"]) <= TOL_COST * g["final_cost"]
    # (the persistent kernel runs the Chronopoulos-Gear form of the same CG: equal iterates in exact arithmetic, the 1e-10 exit can move by one)
    assert abs(out["pcg_iters_history"][1] - g["pcg_iters_history"][1]) <= 3
    for a, b in zip(out["cost_history"], g["cost_history"]):
        assert abs(a - b) <= TOL_COST * max(b, 1.0)


def test_lm_against_oracle_live_and_in_place_update(oracle):
    sc = synthetic.make_scene("small", seed=123)
    p = oracle.Problem(sc)
    so = p.solve(max_iterations=6, num_threads=1, **FIXED_ITERS)
    s, (raxis, t, pts) = gpu_solver(sc)
    out = s.solve(max_iterations=6, **FIXED_ITERS)
    assert abs(out["final_cost"] - so["final_cost"]) <= TOL_COST * so["final_cost"]
    # parameter blocks updated in place, as ceres::Solve does
    assert relerr(raxis, p.cam_r.reshape(-1, 3)) < 1e-8
    assert relerr(t, p.cam_t.reshape(-1, 3)) < 1e-8
    assert relerr(pts, p.pts.reshape(-1, 3)) < 1e-7
    assert np.array_equal(raxis[0], sc["raxis"][0])      # the fixed camera did not move


def test_convergence_and_termination(oracle):
    sc = synthetic.make_scene("small")
    so = oracle.Problem(sc).solve(max_iterations=50, num_threads=1)
    s, _ = gpu_solver(sc)
    out = s.solve(max_iterations=50)
    assert out["status"] == 1 and so["status"] == 1
    assert out["iterations"] == so["iterations"]
    assert abs(out["final_cost"] - so["final_cost"]) <= TOL_COST * so["final_cost"]


def test_deterministic_rerun():
    sc = synthetic.make_scene("small")
    costs = []
    for _ in range(2):
        s, _ = gpu_solver(sc)
        costs.append(s.solve(max_iterations=4, **FIXED_ITERS)["cost_history"])
        s.close()
    assert costs[0] == costs[1]          # bitwise: no floating-point atomics anywhere


def test_pcg_batched_check_same_result():
    sc = synthetic.make_scene("small")
    s1, _ = gpu_solver(sc)
    a = s1.solve(max_iterations=3, **FIXED_ITERS)
    s2, _ = gpu_solver(sc, pcg_check_every=8)
    b = s2.solve(max_iterations=3, **FIXED_ITERS)
    assert a["cost_history"] == b["cost_history"] and a["pcg_iters_history"] == b["pcg_iters_history"]


def test_triangulation(oracle):
    sc = synthetic.make_scene("small")
    sc["raxis"], sc["t"] = sc["raxis_true"].copy(), sc["t_true"].copy()
    order = np.argsort(sc["obs_pt"], kind="stable")
    view_cam = sc["obs_cam"][order]; view_uv = sc["obs_uv"][order]
    view_off = np.concatenate([[0], np.cumsum(np.bincount(sc["obs_pt"], minlength=sc["n_pts"]))]).astype(np.int32)
    xo, oko = oracle.Problem(sc).triangulate(view_off, view_cam, view_uv)
    s, _ = gpu_solver(sc)
    x, ok, ms = s.triangulate(view_off, view_cam, view_uv)
    assert ok.all() and oko.all()
    assert relerr(x, xo) < 1e-10
    assert np.abs(x - sc["pts_true"]).max() < 0.5         # noisy pixels, sane geometry


def _add_heavy_points(sc, pts_idx):
    """Makes the given points visible from EVERY camera (k = n_cams > 32: the 'heavy' path of the point-order kernels)."""
    R = synthetic.axis_angle_to_rotmat(sc["raxis_true"])
    cams, ptsl, uvs = [], [], []
    for p in pts_idx:
        have = set(sc["obs_cam"][sc["obs_pt"] == p].tolist())
        for c in range(sc["n_cams"]):
            if c in have:
                continue
            Pb = R[c].T @ (sc["pts_true"][p] - sc["t_true"][c])
            assert Pb[2] > 0
            uv = synthetic.uv_from_normalized((Pb[:2] / Pb[2])[None, :], sc["intrinsics"][0])[0]
            cams.append(c); ptsl.append(p); uvs.append(uv)
    sc["obs_cam"] = np.ascontiguousarray(np.concatenate([sc["obs_cam"], np.array(cams, np.int32)]))
    sc["obs_pt"] = np.ascontiguousarray(np.concatenate([sc["obs_pt"], np.array(ptsl, np.int32)]))
    sc["obs_uv"] = np.ascontiguousarray(np.concatenate([sc["obs_uv"], np.array(uvs)]))
    sc["n_obs"] = sc["obs_cam"].size
    return sc


def test_heavy_points_more_than_32_views(oracle):
    sc = synthetic.make_scene(None, n_cams=70, n_pts=300, n_obs=1800, seed=5)
    sc = _add_heavy_points(sc, [0, 7, 299])          # 70 views each: three windows per point
    assert np.bincount(sc["obs_pt"]).max() == 70
    p = oracle.Problem(sc)
    o = p.normal_blocks()
    s, _ = gpu_solver(sc)
    g = s.normal_blocks()
    for k in ("gc", "gp", "Hcc", "Hpp"):
        assert relerr(g[k], o[k]) < 1e-11, k
    rng = np.random.default_rng(1)
    lam_c = rng.uniform(0.5, 2.0, (sc["n_cams"], 6)); lam_p = rng.uniform(0.5, 2.0, (sc["n_pts"], 3))
    x = rng.normal(size=(sc["n_cams"], 6))
    yo, ro = p.schur_matvec(lam_c, lam_p, x)
    y, r = s.schur_matvec(lam_c, lam_p, x)
    assert relerr(y, yo) < 1e-11 and relerr(r, ro) < 1e-11
    so = oracle.Problem(sc).solve(max_iterations=4, num_threads=1, **FIXED_ITERS)
    out = s.solve(max_iterations=4, **FIXED_ITERS)
    assert abs(out["final_cost"] - so["final_cost"]) <= TOL_COST * so["final_cost"]


def test_empty_and_ragged():
    # a camera with no observation, a point with no observation, priors only on a point
    sc = synthetic.make_scene("tiny")
    keep = sc["obs_cam"] != 3
    for k in ("obs_cam", "obs_pt", "obs_uv"):
        sc[k] = np.ascontiguousarray(sc[k][keep])
    sc["n_obs"] = int(keep.sum())
    s, _ = gpu_solver(sc)
    e = s.evaluate()
    assert e["ok"] and e["residuals"].shape[0] == sc["n_obs"]
    out = s.solve(max_iterations=3, **FIXED_ITERS)
    assert np.isfinite(out["final_cost"])


def test_venice_scale_properties():
    """Full BASELINE size (1778 / 993923 / 5001946): size-independent properties instead of the oracle."""
    sc = synthetic.make_scene("venice")
    s, _ = gpu_solver(sc, pcg_check_every=4)
    s.finalize()
    assert s.sizes()["n_obs"] == 5001946
    e = s.evaluate(residuals=True, jacobians=False)
    # cost equals half the squared norm of the returned residuals (+ the 4 GCP priors)
    pr = ((sc["pts"][sc["prior_index"]] - sc["prior_b"]) / 0.01) ** 2
    assert abs(e["cost"] - 0.5 * ((e["residuals"] ** 2).sum() + pr.sum())) <= 1e-10 * e["cost"]
    out = s.solve(max_iterations=3, **FIXED_ITERS)
    h = out["cost_history"]
    assert h[1] < h[0] and h[-1] <= h[1]                  # monotone decrease on accepted steps
    # at the truth the RMS residual is the pixel noise of the generator
    # RMS residual at the optimum ~ sigma * sqrt(1 - n_params / n_residuals) = 0.5 * sqrt(1 - 2.99e6 / 1.0e7) = 0.418
    assert 0.38 < np.sqrt(2 * h[-1] / (2 * sc["n_obs"])) < 0.46


@pytest.mark.parametrize("shape", ["small", "ladybug"])
def test_persistent_pcg_kernel_matches_two_kernel_path(shape, monkeypatch):
    """k_pcg_solve (one cooperative launch per LM iteration, TMA ring) against the mat-vec + fused-kernel path on the same
    explicit tiles: same CG (only the summation order of p.q differs), so the iteration counts agree and the costs agree far
    inside the 1e-8 gate; the persistent kernel is the one that runs by default."""
    sc = synthetic.make_scene(shape)
    out = {}
    for persistent in ("1", "0"):
        monkeypatch.setenv("SBA_PCG_PERSISTENT", persistent)
        s, _ = gpu_solver(sc, operator_mode=3)
        s.finalize()
        assert s.explicit_shard()["persistent_pcg"] == int(persistent)
        out[persistent] = s.solve(max_iterations=4, **FIXED_ITERS)
        s.close()
    a, b = out["1"], out["0"]
    # persistent kernel: Chronopoulos-Gear recurrence (one grid barrier per iteration); two-kernel path: the classic loop
    assert all(abs(x - y) <= max(4, 0.08 * y) for x, y in zip(a["pcg_iters_history"], b["pcg_iters_history"])), (a["pcg_iters_history"], b["pcg_iters_history"])
    for x, y in zip(a["cost_history"], b["cost_history"]):
        assert abs(x - y) <= 1e-9 * max(y, 1.0)
    assert a["kernel_launches"] < b["kernel_launches"]


@pytest.mark.parametrize("case", ["ladybug", "ring300", "sparse500"])
def test_block_sparse_panels_match_dense_tiles(case, monkeypatch):
    """k_pcg_solve on the block-sparse panels (only the 6x6 blocks of camera pairs that share a point, compacted from the
    dense tiles after every build; byte ring of variable-size chunks) against the same kernel on the dense tiles: same
    system, only the summation order inside a row / column differs -> same CG iteration counts (+-1 at the 1e-10 exit) and
    costs far inside the 1e-8 gate.  Cases: one partly filled tile; several tiles with a band structure (cameras see their
    neighbours); many tiles, most of them nearly empty (hundreds of tiny panels per chunk)."""
    if case == "ladybug":
        sc = synthetic.make_scene("ladybug")
    elif case == "ring300":
        sc = synthetic.make_scene(None, n_cams=300, n_pts=30000, n_obs=150000, seed=7)
    else:
        sc = synthetic.make_scene(None, n_cams=500, n_pts=12000, n_obs=40000, seed=11)
    out = {}
    for sparse in ("1", "0"):
        monkeypatch.setenv("SBA_PS_SPARSE", sparse)
        s, _ = gpu_solver(sc, operator_mode=3)
        s.finalize()
        assert s.explicit_shard()["persistent_pcg"] == 1
        info = s.explicit_sparse_info()
        assert info["enabled"] == int(sparse)
        if sparse == "1":
            assert info["n_blocks"] >= sc["n_cams"] and info["n_chunks"] <= info["n_panels"]
            assert info["stream_bytes"] >= 288 * info["n_blocks"]
        out[sparse] = s.solve(max_iterations=4, **FIXED_ITERS)
        s.close()
    a, b = out["1"], out["0"]
    # (late LM iterations solve for a step at rounding level: the 1e-10 exit of their CG moves by a few iterations with the summation order)
    assert all(abs(x - y) <= max(4, 0.08 * y) for x, y in zip(a["pcg_iters_history"], b["pcg_iters_history"])), (a["pcg_iters_history"], b["pcg_iters_history"])
    for x, y in zip(a["cost_history"], b["cost_history"]):
        assert abs(x - y) <= 1e-10 * max(y, 1.0)


def test_direct_panel_assembly_equals_the_two_pass_build(monkeypatch):
    """Direct mode (the assembly writes the panels itself, SBA_EX_DIRECT, default on with the panels on one GPU) against the
    dense-tile build + compaction: bitwise equal LM, and the dense tiles the hooks read (refreshed on demand from the panels)
    equal to the ones the two-pass build wrote."""
    sc = synthetic.make_scene(None, n_cams=300, n_pts=30000, n_obs=150000, seed=7)
    monkeypatch.setenv("SBA_PS_SPARSE", "1")
    out, tiles, mv = {}, {}, {}
    rng = np.random.default_rng(5)
    x = rng.normal(size=(sc["n_cams"], 6)); x[0] = 0.0     # zero at the fixed camera
    lam_c = rng.uniform(0.5, 2.0, (sc["n_cams"], 6)); lam_p = rng.uniform(0.5, 2.0, (sc["n_pts"], 3))
    for direct in ("1", "0"):
        monkeypatch.setenv("SBA_EX_DIRECT", direct)
        s, _ = gpu_solver(sc, operator_mode=3)
        out[direct] = s.solve(max_iterations=3, **FIXED_ITERS)
        tiles[direct] = s.explicit_dense()
        mv[direct] = s.schur_matvec(lam_c, lam_p, x)      # its own build with these diagonals, then the tile mat-vec
        s.close()
    assert out["1"]["cost_history"] == out["0"]["cost_history"] and out["1"]["pcg_iters_history"] == out["0"]["pcg_iters_history"]
    assert np.array_equal(tiles["1"][0], tiles["0"][0]) and np.array_equal(tiles["1"][1], tiles["0"][1])
    assert np.array_equal(mv["1"][0], mv["0"][0]) and np.array_equal(mv["1"][1], mv["0"][1])


def test_persistent_pcg_iteration_cap_and_rerun():
    """The kernel stops at the iteration cap like the sequential loop, and is bitwise reproducible."""
    sc = synthetic.make_scene("ladybug")
    hist = []
    for _ in range(2):
        s, _ = gpu_solver(sc, operator_mode=3, pcg_max_iters=7)
        r = s.solve(max_iterations=2, **FIXED_ITERS)
        assert r["pcg_iters_history"][1:] == [7, 7]
        hist.append(r["cost_history"])
        s.close()
    assert hist[0] == hist[1]


def test_two_gpu_sharded_solve_matches_single_gpu():
    """Points sharded over 2 GPUs (torchrun, one process per GPU): explicit tiles reduce-scattered, the per-iteration all-reduce
    inside the persistent kernel over peer mailboxes, and the NCCL fallbacks -- all within 1e-8 of the 1-GPU cost history, with
    bit-identical replicated cameras (scripts/multigpu_check.py asserts both)."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(root, "scripts", "multigpu_check.py"), "ladybug", "3"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=280)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("{")][-1]
    res = json.loads(line)
    assert res["ok"] and {m["mode"] for m in res["modes"]} == {"explicit+peer", "explicit+nccl", "implicit+peer", "implicit+nccl"}
    assert res["modes"][0]["info"][0]["peer_exchange"] == 1 and res["modes"][0]["info"][0]["persistent_pcg"] == 1


def test_two_gpu_block_sparse_panels():
    """2 GPUs, several tiles per rank, block-sparse panels forced on: the block structure is the union over the ranks (a rank
    compacts the tiles it owns, whose blocks come from every rank's points); same 1e-8 / bit-identical-cameras gates."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29534", os.path.join(root, "scripts", "multigpu_check.py"), "cams300", "3", "explicit+peer"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=280, env=dict(os.environ, SBA_PS_SPARSE="1"))
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    res = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][-1])
    assert res["ok"] and all(i["sparse"] == 1 and i["persistent_pcg"] == 1 for i in res["modes"][0]["info"])


def test_two_gpu_shared_blocks_and_pair_factors():
    """2 GPUs with free intrinsics, sensor offsets, free lever arms and UV <-> UV factors: the border / Schur terms of the shared
    blocks are partial sums over a rank's points (one all-reduce), the factor terms are replicated and added once."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29535", os.path.join(root, "scripts", "multigpu_check.py"), "ladybug", "4", "all", "ext"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=280)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    res = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][-1])
    assert res["ok"] and {m["mode"] for m in res["modes"]} == {"explicit+peer", "explicit+nccl"}


def test_venice_scale_explicit_tiles_against_implicit_operator(monkeypatch):
    """Full BASELINE size: the explicit reduced camera system (406 dense tiles built from 11.6 M point-pair contributions) and
    the implicit operator (two passes over the 5 M observations) are independent code paths for the same S; they must agree
    on S x and on the reduced right-hand side.  Then the persistent PCG kernel against the two-kernel path on those tiles."""
    sc = synthetic.make_scene("venice")
    rng = np.random.default_rng(7)
    lam_c = rng.uniform(0.5, 2.0, (sc["n_cams"], 6)); lam_p = rng.uniform(0.5, 2.0, (sc["n_pts"], 3))
    x = rng.normal(size=(sc["n_cams"], 6))
    res = {}
    for mode in (2, 3):
        s, _ = gpu_solver(sc, operator_mode=mode)
        res[mode] = s.schur_matvec(lam_c, lam_p, x)
        if mode == 3:
            info = s.explicit_info()
            assert info["enabled"] and info["n_tiles"] == 406
        s.close()
    assert relerr(res[3][0], res[2][0]) < 1e-11 and relerr(res[3][1], res[2][1]) < 1e-11
    out = {}
    for persistent in ("1", "0"):
        monkeypatch.setenv("SBA_PCG_PERSISTENT", persistent)
        s, _ = gpu_solver(sc, pcg_check_every=8)
        out[persistent] = s.solve(max_iterations=2, **FIXED_ITERS)
        s.close()
    for a, b in zip(out["1"]["cost_history"], out["0"]["cost_history"]):
        assert abs(a - b) <= 1e-9 * b
    assert all(abs(a - b) <= max(4, 0.08 * b) for a, b in zip(out["1"]["pcg_iters_history"], out["0"]["pcg_iters_history"]))


@pytest.mark.parametrize("kw", [dict(), dict(mixed=True)])
@pytest.mark.parametrize("with_info", [False, True])
def test_intrinsics_jacobian(oracle, kw, with_info):
    """SURVEY 8(f) row 1 (first half): the CUDA intrinsics-Jacobian hook against the oracle's duals over the five intrinsic blocks."""
    sc = synthetic.make_scene("small", **kw)
    if with_info:
        sc["obs_info"] = np.random.default_rng(5).normal(size=(sc["n_obs"], 2, 2)) + 2 * np.eye(2)
    jo = oracle.Problem(sc).intrinsics_jac()
    s, _ = gpu_solver(sc)
    j = s.intrinsics_jacobian()
    assert j.shape == jo.shape and np.abs(j[:, :, 8:]).max() == 0.0
    for c in range(8):
        assert relerr(j[:, :, c], jo[:, :, c]) < TOL_J, c


def _with_priors(sc, rng):
    """Pose priors on rAxis / t (imagealignementsbamodule.cpp:256-286) and point priors, all with FULL non-symmetric
    stiffness matrices (FixedSizeNormalPrior<3,3>::Evaluate is r = A (x - b), J = A row-major, fixedsizenormalprior.h:75-112);
    two priors on one block, a prior on the fixed camera (skipped in the blocks), a prior on a fixed point."""
    sc = dict(sc)
    kind = [1, 2, 1, 2, 2, 1, 0, 0, 0, 0]
    index = [3, 3, 7, 7, 7, 0, 11, 11, 40, 5]
    A = rng.normal(size=(len(kind), 3, 3)) * 3.0 + 2.0 * np.eye(3)
    b = np.empty((len(kind), 3))
    for i, (k, j) in enumerate(zip(kind, index)):
        x = sc["pts_true"][j] if k == 0 else (sc["raxis_true"][j] if k == 1 else sc["t_true"][j])
        b[i] = x + rng.normal(0, 0.02, 3)
    sc["pt_fixed"] = sc["pt_fixed"].copy(); sc["pt_fixed"][5] = 1
    sc["prior_kind"] = np.concatenate([sc["prior_kind"], np.array(kind, np.int32)])
    sc["prior_index"] = np.concatenate([sc["prior_index"], np.array(index, np.int32)])
    sc["prior_A"] = np.concatenate([sc["prior_A"], A.reshape(-1, 9)])
    sc["prior_b"] = np.concatenate([sc["prior_b"], b])
    return sc


@pytest.mark.parametrize("operator_mode", [0, 2])
def test_pose_priors_and_full_prior_matrices(oracle, operator_mode):
    sc = _with_priors(synthetic.make_scene("small"), np.random.default_rng(17))
    p = oracle.Problem(sc)
    o = p.normal_blocks()
    eo = p.evaluate()
    s, _ = gpu_solver(sc, operator_mode=operator_mode)
    e = s.evaluate(jacobians=False)
    assert abs(e["cost"] - eo["cost"]) <= 1e-12 * eo["cost"]
    g = s.normal_blocks()
    for k in ("gc", "gp", "Hcc", "Hpp"):
        assert relerr(g[k], o[k]) < 1e-11, k
    so = oracle.Problem(sc).solve(max_iterations=5, num_threads=1, **FIXED_ITERS)
    out = s.solve(max_iterations=5, **FIXED_ITERS)
    assert abs(out["pcg_iters_history"][1] - so["pcg_iters_history"][1]) <= 3
    for a, b in zip(out["cost_history"], so["cost_history"]):
        assert abs(a - b) <= TOL_COST * max(b, 1.0)


def test_prior_only_point(oracle):
    """A free point that carries priors but no observation (a ground-control landmark no image sees): its blocks are the
    priors alone -- after ANY number of linearisations -- and the LM moves it onto the prior like the oracle does."""
    sc = synthetic.make_scene("small")
    lonely = 17
    keep = sc["obs_pt"] != lonely
    for k in ("obs_cam", "obs_pt", "obs_uv"):
        sc[k] = np.ascontiguousarray(sc[k][keep])
    sc["n_obs"] = int(keep.sum())
    rng = np.random.default_rng(2)
    A = rng.normal(size=(2, 3, 3)) + 3.0 * np.eye(3)
    sc["prior_kind"] = np.concatenate([sc["prior_kind"], np.zeros(2, np.int32)])
    sc["prior_index"] = np.concatenate([sc["prior_index"], np.full(2, lonely, np.int32)])
    sc["prior_A"] = np.concatenate([sc["prior_A"], A.reshape(-1, 9)])
    sc["prior_b"] = np.concatenate([sc["prior_b"], sc["pts_true"][lonely] + rng.normal(0, 0.01, (2, 3))])
    p = oracle.Problem(sc)
    o = p.normal_blocks()
    s, (_, _, pts) = gpu_solver(sc)
    for _ in range(3):                       # the blocks must not grow with repeated linearisations
        g = s.normal_blocks()
        for k in ("gp", "Hpp"):
            assert relerr(g[k], o[k]) < 1e-11, k
    so = p.solve(max_iterations=3, num_threads=1, **FIXED_ITERS)
    out = s.solve(max_iterations=3, **FIXED_ITERS)
    for a, b in zip(out["cost_history"], so["cost_history"]):
        assert abs(a - b) <= TOL_COST * max(b, 1.0)
    assert np.abs(pts[lonely] - p.pts.reshape(-1, 3)[lonely]).max() < 1e-9
    assert np.abs(pts[lonely] - sc["pts"][lonely]).max() > 1e-3      # it did move


def test_iteration_callback_reports_and_aborts():
    """sba_set_iteration_callback = ModularSBASolverIterationCallback (modularsbasolver.cpp:30-56): called once per LM
    iteration with the state the summary records; a non-zero return stops the solve after that iteration."""
    sc = synthetic.make_scene("small")
    s, _ = gpu_solver(sc)
    seen = []
    s.set_iteration_callback(lambda info: seen.append(dict(info)) and False)
    out = s.solve(max_iterations=5, **FIXED_ITERS)
    assert [i["iteration"] for i in seen] == [1, 2, 3, 4, 5]
    for i in seen:
        k = i["iteration"]
        assert i["cost"] == out["cost_history"][k] and i["radius"] == out["radius_history"][k]
        assert i["pcg_iterations"] == out["pcg_iters_history"][k]
        assert i["step_valid"] == 1
        if i["step_accepted"]:
            assert abs(i["cost_change"] - (out["cost_history"][k - 1] - out["cost_history"][k])) <= 1e-9 * out["cost_history"][k - 1]
    s.close()
    s, _ = gpu_solver(sc)
    calls = []
    s.set_iteration_callback(lambda info: calls.append(info["iteration"]) or info["iteration"] == 2)
    out2 = s.solve(max_iterations=5, **FIXED_ITERS)
    assert calls == [1, 2] and out2["iterations"] == 2
    assert out2["cost_history"] == out["cost_history"][:3]
    s.set_iteration_callback(None)
    out3 = s.solve(max_iterations=2, copy_params=False, **FIXED_ITERS)
    assert calls == [1, 2] and out3["iterations"] == 2


def test_triangulation_mixed_scene_100k_points(oracle):
    """K5 on a mixed plane / spherical scene with > 1e5 points against the oracle's DLT."""
    sc = synthetic.make_scene(None, n_cams=240, n_pts=120_000, n_obs=600_000, seed=21, mixed=True)
    sc["raxis"], sc["t"] = sc["raxis_true"].copy(), sc["t_true"].copy()
    order = np.argsort(sc["obs_pt"], kind="stable")
    view_cam = sc["obs_cam"][order]; view_uv = sc["obs_uv"][order]
    view_off = np.concatenate([[0], np.cumsum(np.bincount(sc["obs_pt"], minlength=sc["n_pts"]))]).astype(np.int32)
    xo, oko = oracle.Problem(sc).triangulate(view_off, view_cam, view_uv)
    s, _ = gpu_solver(sc)
    x, ok, ms = s.triangulate(view_off, view_cam, view_uv)
    assert ok.all() and oko.all()
    assert relerr(x, xo) < 1e-10
    assert np.median(np.abs(x - sc["pts_true"]).max(axis=1)) < 0.05


@pytest.mark.parametrize("mixed", [False, True])
def test_venice_scale_against_oracle(oracle, mixed):
    """The north-star target itself: the full Venice-shape scene (1778 / 993 923 / 5 001 946) on the GPU against the CPU
    oracle on the same arrays -- all residuals and Jacobians at 1e-12, and the cost history of 3 fixed LM iterations at
    1e-8 with equal PCG iteration counts (+-1 at the 1e-10 exit)."""
    oracle.set_max_threads(0)
    sc = synthetic.make_scene("venice", mixed=mixed)
    p = oracle.Problem(sc)
    o = p.evaluate(threads=0)
    s, _ = gpu_solver(sc, pcg_check_every=4)
    e = s.evaluate()
    assert e["ok"] and o["ok"] and e["residuals"].shape == (5001946, 2)
    assert relerr(e["residuals"], o["residuals"]) < TOL_J
    assert relerr(e["jac_cam"], o["jac_cam"]) < TOL_J
    assert relerr(e["jac_pt"], o["jac_pt"]) < TOL_J
    assert abs(e["cost"] - o["cost"]) <= 1e-12 * o["cost"]
    del e, o
    so = p.solve(max_iterations=3, num_threads=0, **FIXED_ITERS)
    out = s.solve(max_iterations=3, **FIXED_ITERS)
    assert out["iterations"] == so["iterations"] == 3
    for a, b in zip(out["cost_history"], so["cost_history"]):
        assert abs(a - b) <= TOL_COST * b
    assert abs(out["final_cost"] - so["final_cost"]) <= TOL_COST * so["final_cost"]
    # same CG on the same system; the panels sum rows / columns in another order than the oracle's implicit operator, so the
    # exit at 1e-10 (on a plateau of the residual) may fall a few iterations apart
    print("PCG iterations, GPU vs oracle:", out["pcg_iters_history"][1:], so["pcg_iters_history"][1:])
    gi, oi = out["pcg_iters_history"][1:], so["pcg_iters_history"][1:]
    assert all(abs(a - b) <= max(4, 0.08 * b) for a, b in zip(gi[:2], oi[:2]))
    assert all(abs(a - b) <= 0.12 * b for a, b in zip(gi[2:], oi[2:]))   # third step: |delta| is at the rounding floor of the 1e-10 exit (130 vs 143 measured)
