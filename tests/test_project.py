"""".steviapproj" ingest / write-back (SURVEY 8f row 4, I/O half) without a GPU: sba_project_* against a plain-Python statement of
the reference's rules (files cited in tests/project_files.py and csrc/sba_project.cpp)."""
import json

import numpy as np
import pytest

import project_files as pf
from steviapp_b200 import project

INTR = [2000.0, 1990.0, 1490.0, -0.05, 0.01, 0.002, 1e-4, -2e-4, 3e-4, -1e-4]


def simple_project(tmp_path, n_img=3, n_lm=6, **kw):
    """n_img images seeing all n_lm landmarks; image 0 carries position + rotation priors."""
    rng = np.random.default_rng(0)
    lms = [pf.landmark(100 + l, opt=rng.normal(0, 5, 3)) for l in range(n_lm)]
    imgs = []
    for i in range(n_img):
        obs = [pf.image_landmark(1000 * (i + 1) + l, 100 + l, rng.uniform(0, 3000, 2)) for l in range(n_lm)]
        imgs.append(pf.image(10 + i, 1, obs, opt_t=rng.normal(0, 20, 3), opt_r=rng.normal(0, 1, 3)))
    cams = [pf.camera(1, INTR)]
    path = str(tmp_path / "p.steviapproj")
    doc = pf.write(path, kw.pop("landmarks", lms), kw.pop("images", imgs), kw.pop("cameras", cams), **kw)
    return path, doc


def test_values_observations_and_order(tmp_path):
    path, doc = simple_project(tmp_path)
    with project.Project(path) as p:
        assert (p.n_models, p.n_images, p.n_landmarks, p.n_obs, p.n_priors) == (1, 3, 6, 18, 0)
        assert list(p.image_id) == [10, 11, 12] and list(p.landmark_id) == list(range(100, 106)) and list(p.model_id) == [1]
        sc = p.scene()
    for i, im in enumerate(doc[pf.IMAGE]):
        assert np.array_equal(sc["t"][i], im["op"]["vals"]) and np.array_equal(sc["raxis"][i], im["or"]["vals"])
        for l, iml in enumerate(im["Landmarks"]):
            k = 6 * i + l
            assert sc["obs_cam"][k] == i and sc["obs_pt"][k] == l
            assert np.array_equal(sc["obs_uv"][k], [iml["x"]["val"], iml["y"]["val"]])
    assert np.array_equal(sc["obs_info"], np.tile([1.0, 0, 0, 1.0], (18, 1)))
    for l, lm in enumerate(doc[pf.LANDMARK]):
        assert np.array_equal(sc["pts"][l], lm["op"]["vals"])
    assert np.array_equal(sc["intrinsics"][0], INTR)
    assert sc["intr_fixed_mask"][0] == 0 and not sc["cam_fixed"].any() and not sc["pt_fixed"].any()
    # FixedParameter::CameraInternal: every block constant, no free-intrinsics keys
    sc2 = project.load_project(path, project.FIX_CAMERA_INTERNAL)
    assert "intr_fixed_mask" not in sc2 and np.array_equal(sc2["intrinsics"][0], INTR)


def test_graph_reduction_and_disabled_items(tmp_path):
    rng = np.random.default_rng(1)
    lms = [pf.landmark(100 + l, opt=rng.normal(0, 5, 3)) for l in range(6)]
    lms.append(pf.landmark(200, opt=[1, 2, 3]))                       # seen once, no prior: 2 < 3 observations -> dropped
    lms.append(pf.landmark(201, prior=[4, 5, 6], sigma=[0.1, 0.2, 0.3]))   # seen once + prior: 5 >= 3 -> kept
    lms.append(pf.landmark(202, opt=[0, 0, 1], enabled=False))         # disabled
    def obs(i, ids):
        return [pf.image_landmark(1000 * i + k, lid, rng.uniform(0, 100, 2)) for k, lid in enumerate(ids)]
    imgs = [pf.image(10, 1, obs(1, [100, 101, 102, 103, 104, 105, 200, 202]), opt_t=[0, 0, 0], opt_r=[0.1, 0, 0]),
            pf.image(11, 1, obs(2, [100, 101, 102, 103, 104, 105, 201]), opt_t=[1, 0, 0], opt_r=[0, 0.1, 0]),
            pf.image(12, 1, obs(3, [100, 101]), opt_t=[2, 0, 0], opt_r=[0, 0, 0.1]),          # 4 < 6 -> dropped
            pf.image(13, 1, obs(4, [100, 101, 102, 103, 104, 105]), opt_t=[3, 0, 0], enabled=False),
            pf.image(14, 99, obs(5, [100, 101, 102]), opt_t=[4, 0, 0])]                        # no such camera
    path, _ = simple_project(tmp_path, landmarks=lms, images=imgs)
    with project.Project(path) as p:
        assert list(p.image_id) == [10, 11] and p.n_images_without_camera == 1
        assert list(p.landmark_id) == [100, 101, 102, 103, 104, 105, 201]
        sc = p.scene()
    assert sc["n_obs"] == 13                      # 6 + 6 + landmark 201; the observations of 200 / 202 are gone
    assert sc["prior_kind"].tolist() == [0] and sc["prior_index"].tolist() == [6]
    assert np.allclose(np.diag(sc["prior_A"][0].reshape(3, 3)), [10.0, 5.0, 1 / 0.3]) and np.array_equal(sc["prior_b"][0], [4, 5, 6])
    assert np.array_equal(sc["pts"][6], [4, 5, 6])   # no optimised value: the prior is the initial value
