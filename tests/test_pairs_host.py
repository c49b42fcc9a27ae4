"""UV <-> UV cross-projection cost without a GPU: the product's functor on its own forward-mode duals
(steviapp_b200/csrc/sba_jet.cuh, compiled for the host by tests/hostmath) against the oracle's restatement of UV2UVCost
(costfunctors/modularuvprojection.h:660-795), and the oracle against the reference's imagePair known-answer vector
(tests/test_sba_solver/main.cpp:335-492: cost < 1e-6 at the exact solution) and the frame / push-broom cross-projector vector
(tests/test_uv_projectors/main.cpp:258-382)."""
import ctypes as C

import numpy as np

from conftest import ptr, relerr
from steviapp_b200 import synthetic


def image_pair_scene():
    """tests/test_sba_solver/main.cpp:376-440: f = 35, pp = (320, 240), img1 at (r = 0, t = (10, 0, 0)), img2 at
    (r = (0, -pi/2, 0), t = (50, 0, 0)), the five points of the PnP test, uv = f * P.xy / P.z + pp in both images."""
    f = 35.0
    intr = np.array([[f, 320.0, 240.0, 0, 0, 0, 0, 0, 0, 0]])
    r = np.array([[0.0, 0.0, 0.0], [0.0, -np.pi / 2, 0.0]]); t = np.array([[10.0, 0.0, 0.0], [50.0, 0.0, 0.0]])
    pts = np.array([[42, 69, 33], [33, 27, 50], [-6, -99, 70], [-49, -28, 12], [-69, -42, 33.0]])
    R = synthetic.axis_angle_to_rotmat(r)
    uv = []
    for X in pts:
        row = []
        for c in range(2):
            P = R[c].T @ (X - t[c])
            row += [P[0] / P[2] * f + 320.0, P[1] / P[2] * f + 240.0]
        uv.append(row)
    return {"raxis": r, "t": t, "pts": np.zeros((1, 3)), "cam_model": np.zeros(2, np.int32), "intrinsics": intr,
            "obs_cam": np.zeros(0, np.int32), "obs_pt": np.zeros(0, np.int32), "obs_uv": np.zeros((0, 2)),
            "pair_cam": np.array([[0, 1]] * 5, np.int32), "pair_uv": np.array(uv),
            "cam_fixed": np.array([1, 0], np.uint8), "pt_fixed": np.ones(1, np.uint8),
            "prior_kind": np.zeros(0, np.int32), "prior_index": np.zeros(0, np.int32), "prior_A": np.zeros((0, 9)), "prior_b": np.zeros((0, 3)),
            "obs_info": None, "cost_kind": np.zeros(2, np.uint8)}


def pair_scene(shape="small", n_pairs=400, seed=3, noise=0.5, outliers=0.0):
    """A synthetic scene plus UV <-> UV factors: pairs of observations of the same (hidden) point from two cameras."""
    sc = synthetic.make_scene(shape)
    rng = np.random.default_rng(seed)
    order = np.argsort(sc["obs_pt"], kind="stable")
    pt = sc["obs_pt"][order]
    first = np.flatnonzero(np.r_[True, pt[1:] != pt[:-1]])
    cnt = np.diff(np.r_[first, pt.size])
    pick = rng.choice(np.flatnonzero(cnt >= 2), size=n_pairs, replace=False)
    a = order[first[pick]]; b = order[first[pick] + 1]
    sc["pair_cam"] = np.stack([sc["obs_cam"][a], sc["obs_cam"][b]], axis=1).astype(np.int32)
    uv = np.concatenate([sc["obs_uv"][a], sc["obs_uv"][b]], axis=1)
    bad = rng.random(n_pairs) < outliers
    uv[bad] += rng.normal(0, 40.0, (int(bad.sum()), 4))
    sc["pair_uv"] = uv
    return sc


def test_oracle_image_pair_known_answer(oracle):
    e = oracle.ExtProblem(image_pair_scene()).evaluate()
    assert abs(e["cost"]) < 1e-6 and np.abs(e["pair_residuals"]).max() < 1e-9


def cross_projector_scene(sba_lib):
    """tests/test_uv_projectors/main.cpp:258-382 (testCrossUVProjector): a frame camera (f = 35, pp at the image centre, at
    (27, -33, 0)) and a push-broom camera (f = 35, pp = 320, sensor width 640, at the origin, zero polynomials) see the point
    (42, 0, 35); the cross-projection cost is < 1e-6.  The push-broom pixel enters through sba_pushbroom_equivalent_uv."""
    f, ppx, ppy, pbpp = 35.0, 320.0, 240.0, 320.0          # the frame camera's pp cancels out of the vector (pixel = offset + pp)
    X = np.array([42.0, 0.0, f]); t_ph = np.array([27.0, -33.0, 0.0])
    uv_ph = (np.array([X[0] - t_ph[0] + ppx, X[1] - t_ph[1] + ppy]) / X[2]) * f        # main.cpp:336-339
    uv_pb = (np.array([[X[0] + pbpp, X[1]]]) / X[2]) * f                                # main.cpp:341-344
    zero6 = np.zeros(6); uv_eq = np.empty((1, 2)); intr_pb = np.empty(10)
    assert sba_lib.sba_pushbroom_equivalent_uv(1, ptr(np.ascontiguousarray(uv_pb)), 640.0, f, pbpp, ptr(zero6), ptr(zero6), ptr(uv_eq), ptr(intr_pb)) == 0
    sc = image_pair_scene()
    sc["intrinsics"] = np.array([[f, ppx, ppy, 0, 0, 0, 0, 0, 0, 0], intr_pb])
    sc["cam_model"] = np.array([0, 1], np.int32)
    sc["raxis"] = np.zeros((2, 3)); sc["t"] = np.array([t_ph, [0.0, 0.0, 0.0]])
    sc["pair_cam"] = np.array([[0, 1]], np.int32)
    sc["pair_uv"] = np.array([[uv_ph[0], uv_ph[1], uv_eq[0, 0], uv_eq[0, 1]]])
    return sc


def test_oracle_cross_projector_known_answer(oracle, sba_lib):
    sc = cross_projector_scene(sba_lib)
    e = oracle.ExtProblem(sc).evaluate()
    assert abs(e["cost"]) < 1e-6 and np.abs(e["pair_residuals"]).max() < 1e-9


def test_product_functor_matches_oracle_duals(hostmath, oracle):
    sc = pair_scene(n_pairs=300)
    rng = np.random.default_rng(1)
    sc["pair_info"] = np.concatenate([rng.normal(0, 0.2, (300, 2, 2)) + np.eye(2), rng.normal(0, 0.2, (300, 2, 2)) + np.eye(2)], axis=1).reshape(300, 8)
    e = oracle.ExtProblem(sc).evaluate()
    c1, c2 = sc["pair_cam"][:, 0], sc["pair_cam"][:, 1]
    n = c1.size
    r1 = np.ascontiguousarray(sc["raxis"][c1]); t1 = np.ascontiguousarray(sc["t"][c1])
    r2 = np.ascontiguousarray(sc["raxis"][c2]); t2 = np.ascontiguousarray(sc["t"][c2])
    intr = np.ascontiguousarray(sc["intrinsics"][0])
    res = np.empty((n, 4)); jac = np.empty((n, 4, 28))
    hostmath.hm_uv2uv(C.c_int32(n), ptr(r1), ptr(t1), ptr(r2), ptr(t2), ptr(intr), ptr(intr), ptr(np.ascontiguousarray(sc["pair_uv"])),
                      ptr(np.ascontiguousarray(sc["pair_info"])), ptr(res), ptr(jac))
    assert relerr(res, e["pair_residuals"]) < 1e-12
    assert relerr(jac[:, :, :12], e["pair_jac"][:, :, :12]) < 1e-12
    for c in range(8):          # intrinsic columns of camera 1's / camera 2's model (the oracle carries 10 each, B columns zero)
        assert relerr(jac[:, :, 12 + c], e["pair_jac"][:, :, 12 + c]) < 1e-12, c
        assert relerr(jac[:, :, 20 + c], e["pair_jac"][:, :, 22 + c]) < 1e-12, c
    assert np.abs(e["pair_jac"][:, :, 20:22]).max() == 0.0 and np.abs(e["pair_jac"][:, :, 30:32]).max() == 0.0


def test_loss_functions_against_closed_forms(hostmath):
    """rho of ceres::HuberLoss / CauchyLoss / ArctanLoss(a) and the Corrector's scaling (rho'' <= 0 for all three: plain sqrt(rho'))."""
    hostmath.hm_loss.argtypes = [C.c_int, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
    for a in (1.0, 2.5):
        for s in (0.0, 0.3, 1.0, 7.0, 400.0):
            rho = np.empty(3); cor = np.empty(3)
            hostmath.hm_loss(1, a, s, ptr(rho), ptr(cor))
            assert np.isclose(rho[0], s if s <= a * a else 2 * a * np.sqrt(s) - a * a)
            hostmath.hm_loss(2, a, s, ptr(rho), ptr(cor))
            assert np.isclose(rho[0], a * a * np.log1p(s / (a * a))) and np.isclose(rho[1], 1.0 / (1.0 + s / (a * a)))
            hostmath.hm_loss(3, a, s, ptr(rho), ptr(cor))
            assert np.isclose(rho[0], a * np.arctan2(s, a)) and np.isclose(rho[1], 1.0 / (1.0 + s * s / (a * a)))
            assert np.isclose(cor[0], np.sqrt(rho[1])) and cor[1] == cor[0] and cor[2] == 0.0
