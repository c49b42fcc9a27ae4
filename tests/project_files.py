"""Writer of ".steviapproj" files for the tests, restating the reference's toJson key names
(libs/datablocks/project.cpp:1219-1237 envelope, floatparameter.cpp:256-266, point3d.cpp:155-166, rigidbody.cpp:344-361,
image.cpp:368-401,715-723, camera.cpp:548-608).  Test infrastructure only."""
import json

LANDMARK = "StereoVisionApp::Landmark"
IMAGE = "StereoVisionApp::Image"
CAMERA = "StereoVisionApp::Camera"


def fp(val=0.0, is_set=True, stddev=None):
    return {"isSet": int(is_set), "val": float(val), "isUncertain": int(stddev is not None), "stddev": 1.0 if stddev is None else float(stddev)}


def unset():
    return fp(0.0, False)


def group3(vals=None):
    if vals is None:
        return {"isSet": 0, "vals": [0.0, 0.0, 0.0], "isUncertain": 0, "CovMatTrig": []}
    return {"isSet": 1, "vals": [float(v) for v in vals], "isUncertain": 0, "CovMatTrig": []}


def envelope(obj, ident, name, fixed=False, enabled=True):
    obj.update({"id": int(ident), "name": name, "fixed": "fixed" if fixed else "unfixed",
                "enabled": "enabled" if enabled else "disabled", "referers": [], "referered": []})
    return obj


def landmark(ident, prior=None, sigma=None, opt=None, fixed=False, enabled=True, geo_crs=None):
    """prior: (x, y, z) or None; sigma: None (certain) or 3 stddevs; opt: optimised position or None."""
    o = {}
    for k, ax in enumerate("xyz"):
        o[ax] = unset() if prior is None else fp(prior[k], True, None if sigma is None else sigma[k])
    o["op"] = group3(opt)
    if geo_crs:
        o["geo_crs"] = geo_crs
    return envelope(o, ident, f"lm{ident}", fixed, enabled)


def image_landmark(ident, landmark_id, uv, sigma=None):
    o = {"attachedLandmarkId": int(landmark_id), "x": fp(uv[0], True, None if sigma is None else sigma[0]),
         "y": fp(uv[1], True, None if sigma is None else sigma[1])}
    return envelope(o, ident, f"iml{ident}")


def image(ident, camera_id, lms, t_prior=None, t_sigma=None, r_prior=None, r_sigma=None, opt_t=None, opt_r=None,
          fixed=False, enabled=True, trajectory=-1, mounting=-1):
    o = {}
    for k, ax in enumerate("xyz"):
        o[ax] = unset() if t_prior is None else fp(t_prior[k], True, None if t_sigma is None else t_sigma[k])
        o["r" + ax] = unset() if r_prior is None else fp(r_prior[k], True, None if r_sigma is None else r_sigma[k])
    o["op"] = group3(opt_t)
    o["or"] = group3(opt_r)
    o.update({"imFile": f"img{ident}.png", "assignedCamera": int(camera_id), "assignedMounting": int(mounting),
              "assignedTrajectory": int(trajectory), "Landmarks": lms})
    return envelope(o, ident, f"im{ident}", fixed, enabled)


def camera(ident, intr, sigma=None, opt=None, fixed=False, use=(True, True, True), size=(4000, 3000)):
    """intr / opt / sigma: 10 values in the engine's order f cx cy k1 k2 k3 p1 p2 B1 B2 (sigma entries may be None)."""
    keys = ["f", "cx", "cy", "k1", "k2", "k3", "p1", "p2", "B1", "B2"]
    o = {"s": {"w": size[0], "h": size[1]}, "pix_ratio": fp(1.0), "opix_ratio": unset(),
         "useRd": bool(use[0]), "useDRd": False, "useTd": bool(use[1]), "useSd": bool(use[2])}
    for k, key in enumerate(keys):
        o[key] = fp(intr[k], True, None if sigma is None else sigma[k])
        o["o" + key] = unset() if opt is None else fp(opt[k], True)
    for key in ("k4", "k5", "k6"):
        o[key] = fp(0.0)
        o["o" + key] = unset()
    return envelope(o, ident, f"cam{ident}", fixed)


def write(path, landmarks, images, cameras, local_transform=None, extra=None):
    doc = {LANDMARK: landmarks, IMAGE: images, CAMERA: cameras, "metadata": {}}
    if local_transform is not None:
        R, t = local_transform
        doc["metadata"]["local_transform"] = {"R": [[float(v) for v in row] for row in R], "t": [float(v) for v in t]}
    if extra:
        doc.update(extra)
    with open(path, "w") as f:
        json.dump(doc, f, indent=4)
    return doc


def from_scene(path, sc, intr_sigma=None, obs_sigma=None, gcp_sigma=0.01, n_gcp=4, pose_sigma=None):
    """A project file that states a synthetic scene (steviapp_b200.synthetic.make_scene): one Camera per intrinsics model,
    images with their initial poses as optimised values (the first one fixed, its pose as priors), landmarks with their
    initial positions as optimised values, the first n_gcp of them with the true position as an uncertain prior."""
    import numpy as np
    cams = [camera(1 + m, list(sc["intrinsics"][m]), sigma=intr_sigma) for m in range(sc["intrinsics"].shape[0])]
    lms = []
    for l in range(sc["n_pts"]):
        gcp = l < n_gcp
        lms.append(landmark(100000 + l, prior=sc["pts_true"][l] if gcp else None, sigma=[gcp_sigma] * 3 if gcp else None, opt=sc["pts"][l]))
    order = np.argsort(sc["obs_cam"], kind="stable")
    first = np.searchsorted(sc["obs_cam"][order], np.arange(sc["n_cams"] + 1))
    imgs = []
    for c in range(sc["n_cams"]):
        sel = order[first[c]:first[c + 1]]
        obs = [image_landmark(10_000_000 + int(o), 100000 + int(sc["obs_pt"][o]), sc["obs_uv"][o], obs_sigma) for o in sel]
        kw = {}
        if c == 0:
            kw = dict(t_prior=sc["t"][0], r_prior=sc["raxis"][0], fixed=True)
        elif pose_sigma is not None:
            kw = dict(t_prior=sc["t_true"][c], t_sigma=[pose_sigma[1]] * 3, r_prior=sc["raxis_true"][c], r_sigma=[pose_sigma[0]] * 3)
        imgs.append(image(1000 + c, 1 + int(sc["cam_model"][c]), obs, opt_t=sc["t"][c], opt_r=sc["raxis"][c], **kw))
    return write(path, lms, imgs, cams), order
