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


def test_pose_and_landmark_rules(tmp_path):
    rng = np.random.default_rng(2)
    lms = [pf.landmark(100, prior=[1, 2, 3], opt=[1.5, 2.5, 3.5], fixed=True),          # fixed: constant, no prior
           pf.landmark(101, prior=[4, 5, 6], opt=[4.1, 5.1, 6.1]),                       # certain prior: stiffness 1e6
           pf.landmark(102, prior=[7, 8, 9], sigma=[-0.5, 2.0, 4.0]),                    # |stddev|
           pf.landmark(103), pf.landmark(104, opt=[0, 1, 0]), pf.landmark(105, opt=[0, 0, 1])]
    def obs(i):
        return [pf.image_landmark(1000 * i + l, 100 + l, rng.uniform(0, 100, 2), sigma=(0.5, 2.0) if l == 0 else None) for l in range(6)]
    imgs = [pf.image(10, 1, obs(1), t_prior=[1, 1, 1], r_prior=[0.1, 0.2, 0.3], opt_t=[9, 9, 9], opt_r=[0.9, 0.9, 0.9], fixed=True),
            pf.image(11, 1, obs(2), t_prior=[2, 2, 2], t_sigma=[0.1, 0.2, -0.4], r_prior=[0.3, 0.2, 0.1], r_sigma=[0.01, 0.02, 0.04],
                     opt_t=[2.5, 2.5, 2.5]),
            pf.image(12, 1, obs(3), t_prior=[3, 3, 3], opt_r=[0.5, 0.5, 0.5])]
    path, _ = simple_project(tmp_path, landmarks=lms, images=imgs)
    sc = project.load_project(path)
    # modularsbasolver.cpp:289-302: a fixed image takes its priors; otherwise the optimised value when set, else the prior
    assert np.array_equal(sc["t"], [[1, 1, 1], [2.5, 2.5, 2.5], [3, 3, 3]])
    assert np.array_equal(sc["raxis"], [[0.1, 0.2, 0.3], [0.3, 0.2, 0.1], [0.5, 0.5, 0.5]])
    assert sc["cam_fixed"].tolist() == [1, 0, 0] and sc["pt_fixed"].tolist() == [1, 0, 0, 0, 0, 0]
    assert np.array_equal(sc["pts"], [[1.5, 2.5, 3.5], [4.1, 5.1, 6.1], [7, 8, 9], [0, 0, 0], [0, 1, 0], [0, 0, 1]])
    # priors: landmarks first (file order), then per image position, rotation
    assert sc["prior_kind"].tolist() == [0, 0, 2, 1] and sc["prior_index"].tolist() == [1, 2, 1, 1]
    A = sc["prior_A"].reshape(-1, 3, 3)
    assert np.array_equal(A[0], 1e6 * np.eye(3)) and np.array_equal(sc["prior_b"][0], [4, 5, 6])
    assert np.array_equal(A[1], np.diag([2.0, 0.5, 0.25]))
    assert np.array_equal(A[2], np.diag([1 / 0.1, 1 / 0.2, 1 / 0.4])) and np.array_equal(sc["prior_b"][2], [2, 2, 2])
    assert np.array_equal(A[3], np.diag([1 / 0.01, 1 / 0.02, 1 / 0.04])) and np.array_equal(sc["prior_b"][3], [0.3, 0.2, 0.1])
    # observation stiffness: 1 / stddev where uncertain (imagealignementsbamodule.cpp:424-426)
    info = sc["obs_info"].reshape(-1, 2, 2)
    assert np.array_equal(info[0], np.diag([2.0, 0.5])) and np.array_equal(info[1], np.eye(2))


def test_camera_blocks_and_priors(tmp_path):
    sig = [5.0, 2.0, 3.0, 0.01, 0.02, 0.03, 1e-3, 2e-3, None, None]
    opt = [2010.0, 1995.0, 1495.0, -0.06, 0.02, 0.003, 2e-4, -1e-4, 1e-4, 2e-4]
    cams = [pf.camera(1, INTR, sigma=sig, opt=opt),                       # optimised values win; priors on f pp k t, B certain
            pf.camera(2, INTR, sigma=sig, use=(False, True, False)),      # radial and skew switched off
            pf.camera(3, INTR, sigma=sig, fixed=True)]
    rng = np.random.default_rng(3)
    lms = [pf.landmark(100 + l, opt=rng.normal(0, 5, 3)) for l in range(6)]
    imgs = [pf.image(10 + i, cam, [pf.image_landmark(1000 * i + l, 100 + l, rng.uniform(0, 100, 2)) for l in range(6)], opt_t=[i, 0, 0])
            for i, cam in enumerate([2, 1, 3, 1])]
    path, _ = simple_project(tmp_path, landmarks=lms, images=imgs, cameras=cams)
    with project.Project(path) as p:
        assert list(p.model_id) == [2, 1, 3]          # order of first use by an image
        sc = p.scene()
        sc_fixed = p.scene(project.FIX_CAMERA_INTERNAL)
    assert sc["cam_model"].tolist() == [0, 1, 2, 1]
    assert np.array_equal(sc["intrinsics"][1], opt)
    assert np.array_equal(sc["intrinsics"][0], INTR[:3] + [0, 0, 0] + INTR[6:8] + [0, 0])
    assert sc["intr_fixed_mask"].tolist() == [4 | 16, 0, 0x1f]
    f32 = lambda s: float(np.float32(1.0) / np.float32(s))   # "float sigma; A = 1/sigma" (pinholecameraprojectormodule.cpp:332-336)
    got = {(int(m), int(b)): (A, bb) for m, b, A, bb in zip(sc["intr_prior_model"], sc["intr_prior_block"], sc["intr_prior_A"], sc["intr_prior_b"])}
    assert sorted(got) == [(0, 0), (0, 1), (0, 3), (1, 0), (1, 1), (1, 2), (1, 3)]
    A, b = got[(1, 2)]
    assert np.array_equal(A, np.diag([f32(0.01), f32(0.02), f32(0.03)]).ravel()) and np.array_equal(b, INTR[3:6])   # mean = the prior, not the optimised value
    A, b = got[(0, 1)]
    assert np.array_equal(A[:4], [f32(2.0), 0, 0, f32(3.0)]) and not A[4:].any() and np.array_equal(b, [INTR[1], INTR[2], 0])
    A, b = got[(1, 0)]
    assert A[0] == f32(5.0) and not A[1:].any() and b[0] == INTR[0]
    assert "intr_fixed_mask" not in sc_fixed and np.array_equal(sc_fixed["intrinsics"], sc["intrinsics"])


def test_local_transform_applies_to_landmarks_only(tmp_path):
    c, s = np.cos(0.3), np.sin(0.3)
    R = np.array([[c, -s, 0], [s, c, 0], [0, 0, 1.0]]); t = np.array([10.0, -20.0, 5.0])
    path, doc = simple_project(tmp_path, local_transform=(R, t))
    doc[pf.LANDMARK][0].update(pf.landmark(100, prior=[1, 2, 3], sigma=[1, 1, 1]))   # prior only: initial value and mean both move
    with open(path, "w") as f:
        json.dump(doc, f)
    sc = project.load_project(path)
    for l, lm in enumerate(doc[pf.LANDMARK][1:], start=1):
        assert np.allclose(sc["pts"][l], R @ np.array(lm["op"]["vals"]) + t, rtol=0, atol=1e-12)
    assert np.allclose(sc["pts"][0], R @ [1, 2, 3] + t, atol=1e-12) and np.allclose(sc["prior_b"][0], R @ [1, 2, 3] + t, atol=1e-12)
    assert np.array_equal(sc["t"][0], doc[pf.IMAGE][0]["op"]["vals"])   # "TODO: manage geo transforms for the images" in the reference


def test_write_results_and_save(tmp_path):
    path, doc = simple_project(tmp_path, extra={"StereoVisionApp::Trajectory": [{"id": 77, "name": "träj \"q\"", "blob": [1, 2.5, None, True, {"k": "v\n"}]}]})
    doc[pf.IMAGE][2]["fixed"] = "fixed"
    doc[pf.IMAGE][2].update({ax: pf.fp(1.0) for ax in ("x", "y", "z", "rx", "ry", "rz")})
    with open(path, "w") as f:
        json.dump(doc, f)
    out = str(tmp_path / "out.steviapproj")
    rng = np.random.default_rng(5)
    raxis, t, pts = rng.normal(0, 1, (3, 3)), rng.normal(0, 30, (3, 3)), rng.normal(0, 7, (6, 3))
    intr = np.array([INTR]) * 1.01
    with project.Project(path) as p:
        p.write_results(raxis, t, pts, intr)
        p.save(out)
        sc_after = p.scene()            # the handle sees its own results
    saved = json.load(open(out))
    f32 = lambda a: np.asarray(a, np.float32).astype(np.float64)
    for i in range(2):                 # imagealignementsbamodule.cpp:469-483: through float
        assert np.array_equal(saved[pf.IMAGE][i]["op"]["vals"], f32(t[i])) and saved[pf.IMAGE][i]["op"]["isSet"] == 1
        assert np.array_equal(saved[pf.IMAGE][i]["or"]["vals"], f32(raxis[i]))
    assert saved[pf.IMAGE][2]["op"] == doc[pf.IMAGE][2]["op"] and saved[pf.IMAGE][2]["or"] == doc[pf.IMAGE][2]["or"]   # fixed image: skipped (:463-465)
    for l in range(6):
        assert np.array_equal(saved[pf.LANDMARK][l]["op"]["vals"], f32(pts[l]))
    cam = saved[pf.CAMERA][0]
    for k, key in enumerate(["of", "ocx", "ocy", "ok1", "ok2", "ok3", "op1", "op2", "oB1", "oB2"]):
        assert cam[key]["isSet"] == 1 and cam[key]["val"] == f32(intr[0, k])
    assert cam["ok4"]["isSet"] == 0 and cam["opix_ratio"]["isSet"] == 0          # Camera::clearOptimized
    # everything the engine does not know survives, keys come out sorted (QJsonObject order)
    assert saved["StereoVisionApp::Trajectory"] == doc["StereoVisionApp::Trajectory"]
    assert saved[pf.IMAGE][0]["Landmarks"] == doc[pf.IMAGE][0]["Landmarks"] and saved[pf.CAMERA][0]["f"] == doc[pf.CAMERA][0]["f"]
    text = open(out).read()
    top = [ln.split('"')[1] for ln in text.splitlines() if ln.startswith('    "')]
    assert top == sorted(top)
    # a saved file reads back as the float-rounded solution
    sc = project.load_project(out)
    assert np.array_equal(sc["t"][:2], f32(t[:2])) and np.array_equal(sc["pts"], f32(pts)) and np.array_equal(sc["intrinsics"], f32(intr))
    assert np.array_equal(sc["t"][2], [1, 1, 1])
    for key in ("t", "raxis", "pts", "intrinsics"):
        assert np.array_equal(sc[key], sc_after[key])


def test_number_formatting_round_trips(tmp_path):
    vals = [0.1, 1e-300, 123456789.123456789, -2.5e17, 3.0, 1 / 3, 5e-324, 1.7976931348623157e308]
    path, _ = simple_project(tmp_path, extra={"numbers": vals})
    out = str(tmp_path / "n.steviapproj")
    with project.Project(path) as p:
        p.save(out)
    assert json.load(open(out))["numbers"] == vals


@pytest.mark.parametrize("case", ["missing", "syntax", "array", "geo", "trajectory", "trailing"])
def test_errors(tmp_path, case):
    path, doc = simple_project(tmp_path)
    if case == "missing":
        path = str(tmp_path / "nope.steviapproj")
    elif case == "syntax":
        open(path, "w").write(open(path).read()[:-40])
    elif case == "array":
        open(path, "w").write("[1, 2]")
    elif case == "trailing":
        open(path, "a").write("{}")
    elif case == "geo":
        doc[pf.LANDMARK][1]["geo_crs"] = "EPSG:4326"
        json.dump(doc, open(path, "w"))
    elif case == "trajectory":
        doc[pf.IMAGE][1].update({"assignedTrajectory": 5, "assignedMounting": 6})
        json.dump(doc, open(path, "w"))
    with pytest.raises(ValueError) as e:
        project.Project(path)
    msg = str(e.value)
    assert {"missing": "cannot open", "syntax": "JSON error", "array": "not a JSON object", "geo": "geographic CRS",
            "trajectory": "trajectory", "trailing": "trailing"}[case] in msg


def test_empty_project(tmp_path):
    path = str(tmp_path / "e.steviapproj")
    open(path, "w").write("{}")
    with project.Project(path) as p:
        assert (p.n_models, p.n_images, p.n_landmarks, p.n_obs, p.n_priors) == (0, 0, 0, 0, 0)
        sc = p.scene()
    assert sc["pts"].shape == (0, 3) and sc["obs_uv"].shape == (0, 2)
