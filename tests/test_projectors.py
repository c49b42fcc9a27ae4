"""Push-broom observations on the pinhole kernels (sba_pushbroom_equivalent_uv): with constant intrinsics the
cost sees a projector only through its view ray, so pixels rewritten to (d_x, d_y) with the record {f, 0, ...} must give the
SAME ray through the engine's pinhole dirFromUV as PinholePushbroomUVProjector::dirFromUV
(sbamodules/pinholecameraprojectormodule.h:79-116) gives for the original pixels.  CPU only (host code + oracle)."""
import ctypes as C

import numpy as np
import pytest

from conftest import ptr


def test_pushbroom_rays_through_the_pinhole_record(sba_lib, oracle, hostmath):
    rng = np.random.default_rng(4)
    n, W, f, pp = 500, 4096.0, 5300.0, 2051.25
    a6 = np.array([0.3, -1.2, 0.8, 2.5, -0.7, 1.1]); b6 = np.array([-0.1, 0.6, -1.9, 0.4, 3.0, -2.2])
    uv = np.column_stack([rng.uniform(0, W, n), rng.uniform(-2.0, 2.0, n)])
    uv_eq = np.empty((n, 2)); intr = np.empty(10)
    assert sba_lib.sba_pushbroom_equivalent_uv(n, ptr(uv), W, f, pp, ptr(a6), ptr(b6), ptr(uv_eq), ptr(intr)) == 0
    assert intr[0] == f and not intr[1:].any()
    lib = oracle.load()
    lib.oracle_pushbroom_dir.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.oracle_pushbroom_dir.restype = None
    worst = 0.0
    for i in range(n):
        d_ref = np.empty(3); d_eng = np.empty(3)
        lib.oracle_pushbroom_dir(ptr(np.ascontiguousarray(uv[i])), W, f, pp, ptr(a6), ptr(b6), ptr(d_ref))
        lib.oracle_dir_from_uv(ptr(np.ascontiguousarray(uv_eq[i])), ptr(intr), ptr(d_eng))     # the engine's pinhole model (oracle restatement)
        worst = max(worst, np.abs(d_eng - d_ref).max() / max(np.abs(d_ref).max(), 1.0))
    assert worst < 1e-15, worst
    # and an independent numpy statement of the reference formula
    s = (uv[:, 0] - W / 2) / W
    P = np.vander(s, 6, increasing=True)
    assert np.abs(uv_eq[:, 0] - (uv[:, 0] - pp + P @ a6)).max() < 1e-11 and np.abs(uv_eq[:, 1] - (uv[:, 1] + P @ b6)).max() < 1e-12


def test_pushbroom_argument_checks(sba_lib):
    a6 = np.zeros(6); uv = np.zeros((1, 2)); out = np.zeros((1, 2))
    assert sba_lib.sba_pushbroom_equivalent_uv(1, ptr(uv), 0.0, 100.0, 0.0, ptr(a6), ptr(a6), ptr(out), None) != 0   # sensor width
    assert sba_lib.sba_pushbroom_equivalent_uv(1, ptr(uv), 10.0, 0.0, 0.0, ptr(a6), ptr(a6), ptr(out), None) != 0    # focal length
    assert sba_lib.sba_pushbroom_equivalent_uv(1, None, 10.0, 100.0, 0.0, ptr(a6), ptr(a6), ptr(out), None) != 0
    assert sba_lib.sba_pushbroom_equivalent_uv(0, None, 10.0, 100.0, 0.0, ptr(a6), ptr(a6), None, None) == 0
