"""A ".steviapproj" file through the whole engine on the GPU: load (sba_project_*), solve with the reference's default
configuration (free intrinsic blocks with their priors, pinholecameraprojectormodule.cpp:302-533) against the extended
oracle on the same arrays, write the results back the way the modules do, and read them again."""
import json

import numpy as np
import pytest

import project_files as pf
from steviapp_b200 import SBASolver, project, synthetic

pytestmark = pytest.mark.gpu

FIXED_ITERS = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)


@pytest.mark.parametrize("fixed_flags", [0, project.FIX_CAMERA_INTERNAL])
def test_project_solve_and_write_back(tmp_path, oracle, fixed_flags):
    src = synthetic.make_scene("small")
    rng = np.random.default_rng(11)
    src["intrinsics"] = src["intrinsics"].copy()
    if not fixed_flags:
        src["intrinsics"][0, 0] *= 1.003; src["intrinsics"][0, 1:3] += rng.normal(0, 2.0, 2)
    path = str(tmp_path / "scene.steviapproj")
    pf.from_scene(path, src, intr_sigma=[4.0, 3.0, 3.0, 0.01, 0.01, 0.01, 1e-3, 1e-3, None, None], obs_sigma=(0.8, 1.25))
    with project.Project(path) as p:
        assert (p.n_images, p.n_landmarks, p.n_obs) == (src["n_cams"], src["n_pts"], src["n_obs"])
        sc = p.scene(fixed_flags)
        assert ("intr_fixed_mask" in sc) == (not fixed_flags)
        po = oracle.ExtProblem(sc)
        so = po.solve(max_iterations=5, num_threads=1, **FIXED_ITERS)
        s = SBASolver()
        raxis, t, pts = synthetic.load_into(s, sc)
        out = s.solve(max_iterations=5, **FIXED_ITERS)
        assert out["cost_history"][-1] < 1e-2 * out["cost_history"][0]
        for a, b in zip(out["cost_history"], so["cost_history"]):
            assert abs(a - b) <= 1e-8 * max(b, 1.0), (out["cost_history"], so["cost_history"])
        intr = s.intrinsics if not fixed_flags else None
        p.write_results(raxis, t, pts, intr)
        saved = str(tmp_path / "solved.steviapproj")
        p.save(saved)
        s.close()
    doc = json.load(open(saved))
    f32 = lambda a: np.asarray(a, np.float32).astype(np.float64)
    assert np.array_equal(doc[pf.IMAGE][3]["op"]["vals"], f32(t[3])) and np.array_equal(doc[pf.LANDMARK][7]["op"]["vals"], f32(pts[7]))
    assert doc[pf.IMAGE][0]["op"]["vals"] == list(src["t"][0])        # the fixed image keeps what it had
    if not fixed_flags:
        assert doc[pf.CAMERA][0]["of"]["val"] == f32(intr[0, 0]) and abs(intr[0, 0] - 2000.0) < abs(src["intrinsics"][0, 0] - 2000.0)
    back = project.load_project(saved, fixed_flags)
    assert np.array_equal(back["pts"], f32(pts)) and np.array_equal(back["t"][1:], f32(t[1:]))
