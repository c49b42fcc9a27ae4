"""CUDA hook for the constant-sensor-offset cost (sba_evaluate_offset) against the oracle's duals over the
reference's PoseTransform<UVCost, SourceToInitial> functor, 1e-12 (residuals 2e-12: see tests/test_hostmath.py)."""
import numpy as np
import pytest

from conftest import ptr, relerr
from steviapp_b200 import SBASolver, synthetic

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("spherical", [False, True])
@pytest.mark.parametrize("with_info", [False, True])
def test_offset_hook_matches_oracle(oracle, spherical, with_info):
    sc = synthetic.make_scene("small", spherical=spherical)
    rng = np.random.default_rng(23)
    nc, n = sc["n_cams"], sc["n_obs"]
    if with_info:
        sc["obs_info"] = rng.normal(size=(n, 2, 2)) + 2 * np.eye(2)
    ax = rng.normal(size=(nc, 3)); ax /= np.linalg.norm(ax, axis=1)[:, None]
    r_off = ax * rng.uniform(0.0, 0.05, (nc, 1)); t_off = rng.normal(0, 0.05, (nc, 3))   # small: the points stay in front
    r_off[0] = 0.0; t_off[0] = 0.0
    off12 = np.concatenate([np.transpose(synthetic.axis_angle_to_rotmat(r_off), (0, 2, 1)).reshape(nc, 9), t_off], axis=1)
    with SBASolver() as s:
        synthetic.load_into(s, sc)
        plain = s.evaluate()
        out = s.evaluate_offset(off12)
        zero = s.evaluate_offset(np.concatenate([np.tile(np.eye(3).reshape(1, 9), (nc, 1)), np.zeros((nc, 3))], axis=1))
    cam, pt = sc["obs_cam"], sc["obs_pt"]
    c = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    info = None if sc["obs_info"] is None else c(sc["obs_info"])
    ro = np.empty((n, 2)); jco = np.empty((n, 2, 6)); jpo = np.empty((n, 2, 3)); fo = np.zeros(n, np.uint8)
    bad = oracle.load().oracle_eval_offset(n, ptr(c(sc["raxis"][cam])), ptr(c(sc["t"][cam])), ptr(c(r_off[cam])), ptr(c(t_off[cam])),
                                           ptr(c(sc["pts"][pt])), ptr(c(sc["intrinsics"][0])), ptr(c(sc["obs_uv"])), ptr(info),
                                           int(spherical), ptr(ro), ptr(jco), ptr(jpo), ptr(fo))
    assert bad == 0 and out["n_failed"] == 0
    assert relerr(out["residuals"], ro) < 2e-12 and relerr(out["jac_cam"], jco) < 1e-12 and relerr(out["jac_pt"], jpo) < 1e-12
    # identity offsets: the plain evaluation
    assert relerr(zero["residuals"], plain["residuals"]) < 1e-13 and relerr(zero["jac_cam"], plain["jac_cam"]) < 1e-13
    assert relerr(zero["jac_pt"], plain["jac_pt"]) < 1e-13
