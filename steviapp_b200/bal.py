"""BAL text problems through libsba_b200.so (sba_bal_* in include/sba_b200.h).

Format and camera model: the reference's in-tree BAL example, modules/g2o/g2o/examples/bal/bal_example.cpp:168-204
(projection) and :271-340 (file layout).  The parsing and the conversion to the engine's conventions are C++ inside the
library (csrc/sba_bal.cpp); this module only allocates the arrays and returns a scene dict that `synthetic.load_into`
registers on an `SBASolver`.
"""
import ctypes as C
import os
import warnings

import numpy as np

from . import _capi


def _check(rc, what):
    if rc != _capi.SBA_OK:
        raise ValueError(f"{what}: {_capi.load().sba_bal_error().decode()}")


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def read_bal_raw(path):
    """The file's own arrays: dict(n_cams, n_pts, n_obs, cam_idx, pt_idx, xy [n_obs, 2], cam9 [n_cams, 9], xyz [n_pts, 3])."""
    lib = _capi.load()
    h = C.c_void_p()
    _check(lib.sba_bal_open(os.fsencode(path), C.byref(h)), "sba_bal_open")
    try:
        nc, npt, no = C.c_int32(), C.c_int32(), C.c_int32()
        _check(lib.sba_bal_sizes(h, C.byref(nc), C.byref(npt), C.byref(no)), "sba_bal_sizes")
        out = {"n_cams": nc.value, "n_pts": npt.value, "n_obs": no.value,
               "cam_idx": np.empty(no.value, np.int32), "pt_idx": np.empty(no.value, np.int32),
               "xy": np.empty((no.value, 2)), "cam9": np.empty((nc.value, 9)), "xyz": np.empty((npt.value, 3))}
        _check(lib.sba_bal_raw(h, _ptr(out["cam_idx"]), _ptr(out["pt_idx"]), _ptr(out["xy"]), _ptr(out["cam9"]), _ptr(out["xyz"])),
               "sba_bal_raw")
        return out
    finally:
        lib.sba_bal_close(h)


def load_bal(path, fix_first_camera=True):
    """Reads a BAL file and returns a scene dict in the engine's conventions (same keys as synthetic.make_scene):
    one intrinsics model per camera {f, 0, ...}, plane cost, measurements with the file's radial distortion removed.
    Gauge: BAL files carry no priors; by default the first camera is held constant (SetParameterBlockConstant), as the
    reference does for a fixed frame (modularsbasolver.cpp:822-846)."""
    lib = _capi.load()
    h = C.c_void_p()
    _check(lib.sba_bal_open(os.fsencode(path), C.byref(h)), "sba_bal_open")
    try:
        nc, npt, no = C.c_int32(), C.c_int32(), C.c_int32()
        _check(lib.sba_bal_sizes(h, C.byref(nc), C.byref(npt), C.byref(no)), "sba_bal_sizes")
        nc, npt, no = nc.value, npt.value, no.value
        raxis, t = np.empty((nc, 3)), np.empty((nc, 3))
        intr = np.empty((nc, 10))
        model = np.empty(nc, np.int32)
        pts = np.empty((npt, 3))
        oc, op = np.empty(no, np.int32), np.empty(no, np.int32)
        uv = np.empty((no, 2))
        failed = C.c_int32()
        _check(lib.sba_bal_convert(h, _ptr(raxis), _ptr(t), _ptr(intr), _ptr(model), _ptr(pts), _ptr(oc), _ptr(op), _ptr(uv),
                                   C.byref(failed)), "sba_bal_convert")
    finally:
        lib.sba_bal_close(h)
    if failed.value:
        warnings.warn(f"{path}: the radial distortion of {failed.value} of {no} observations could not be removed (no root of the radius "
                      "equation: wide-angle pixels beyond the turning point of the polynomial); they are kept distorted", RuntimeWarning)
    cam_fixed = np.zeros(nc, np.uint8)
    if fix_first_camera:
        cam_fixed[0] = 1
    return {
        "n_cams": nc, "n_pts": npt, "n_obs": no, "undistort_failed": failed.value,
        "intrinsics": intr, "cam_model": model, "cam_fixed": cam_fixed, "cost_kind": np.zeros(nc, np.uint8),
        "pt_fixed": np.zeros(npt, np.uint8), "raxis": raxis, "t": t, "pts": pts,
        "obs_cam": oc, "obs_pt": op, "obs_uv": uv, "obs_info": None,
        "prior_kind": np.zeros(0, np.int32), "prior_index": np.zeros(0, np.int32),
        "prior_A": np.zeros((0, 9)), "prior_b": np.zeros((0, 3)),
    }


def bal_cost(path, raxis=None, t=None, pts=None):
    """Cost and RMS (pixels) of a BAL file in BAL's OWN metric (1/2 sum |projection - xy|^2, bal_example.cpp:168-204), for the
    file's parameters or -- raxis / t / pts given, in the engine's conventions as load_bal returns and a solve updates them --
    for a solution: the figure to quote against published BAL numbers (the engine's own cost is first-order equivalent only)."""
    lib = _capi.load()
    h = C.c_void_p()
    _check(lib.sba_bal_open(os.fsencode(path), C.byref(h)), "sba_bal_open")
    try:
        cam9 = None
        if raxis is not None:
            raxis = np.ascontiguousarray(raxis, np.float64); t = np.ascontiguousarray(t, np.float64)
            cam9 = np.empty((raxis.shape[0], 9))
            _check(lib.sba_bal_cameras_from_engine(h, _ptr(raxis), _ptr(t), _ptr(cam9)), "sba_bal_cameras_from_engine")
        xyz = None if pts is None else np.ascontiguousarray(pts, np.float64)
        cost, rms = C.c_double(), C.c_double()
        _check(lib.sba_bal_cost(h, None if cam9 is None else _ptr(cam9), None if xyz is None else _ptr(xyz), C.byref(cost), C.byref(rms)), "sba_bal_cost")
        return cost.value, rms.value
    finally:
        lib.sba_bal_close(h)


def write_bal(path, cam_idx, pt_idx, xy, cam9, xyz):
    """Writes a BAL file (every double with 17 significant digits: reading it back is bit-exact)."""
    cam_idx = np.ascontiguousarray(cam_idx, np.int32); pt_idx = np.ascontiguousarray(pt_idx, np.int32)
    xy = np.ascontiguousarray(xy, np.float64); cam9 = np.ascontiguousarray(cam9, np.float64); xyz = np.ascontiguousarray(xyz, np.float64)
    _check(_capi.load().sba_bal_write(os.fsencode(path), cam9.shape[0], xyz.shape[0], cam_idx.shape[0],
                                      _ptr(cam_idx), _ptr(pt_idx), _ptr(xy), _ptr(cam9), _ptr(xyz)), "sba_bal_write")
