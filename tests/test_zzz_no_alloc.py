"""sba_triangulate_noalloc (include/sba_b200.h; no counterpart in the reference):
the DLT kernels with every intermediate in one static __device__ array, called through the C ABI on device pointers, against
the existing entry point sba_triangulate on the same inputs -- two calls in a row on different problems, so that anything a
call left in the shared scratch would show.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from conftest import relerr
from steviapp_b200 import SBASolver, _capi, synthetic

pytestmark = pytest.mark.gpu

TOL = 1e-10          # the gate of test_triangulation (tests/test_gpu_parity.py); the two paths run the same kernels


def views_of(sc):
    order = np.argsort(sc["obs_pt"], kind="stable")
    view_off = np.concatenate([[0], np.cumsum(np.bincount(sc["obs_pt"], minlength=sc["n_pts"]))]).astype(np.int32)
    return view_off, np.ascontiguousarray(sc["obs_cam"][order]), np.ascontiguousarray(sc["obs_uv"][order])


def problem(seed, n_cams, n_pts, n_obs, two_models):
    sc = synthetic.make_scene(None, n_cams=n_cams, n_pts=n_pts, n_obs=n_obs, seed=seed, mixed=True)
    sc["raxis"], sc["t"] = sc["raxis_true"].copy(), sc["t_true"].copy()
    if two_models:          # a second intrinsics model with other values: rays (and points) differ, that is fine -- both paths see the same
        intr2 = sc["intrinsics"][0].copy(); intr2[0] *= 1.01; intr2[3] *= 0.5
        sc["intrinsics"] = np.stack([sc["intrinsics"][0], intr2])
        sc["cam_model"] = (np.arange(n_cams) % 2).astype(np.int32)
    return sc


def existing_entry_point(sc, view_off, view_cam, view_uv):
    s = SBASolver()
    synthetic.load_into(s, sc)
    x, ok, _ = s.triangulate(view_off, view_cam, view_uv)
    s.close()
    return x, ok


def no_alloc(torch, sc, view_off, view_cam, view_uv, cam_model=True, want_ok=True):
    """Device tensors in, device tensors out; the entry point itself allocates nothing (the tensors are the caller's)."""
    lib = _capi.load()
    dev = torch.device("cuda:0")
    d = lambda a, dt: torch.as_tensor(np.ascontiguousarray(a), dtype=dt, device=dev)
    raxis, t = d(sc["raxis"], torch.float64), d(sc["t"], torch.float64)
    intr = d(np.asarray(sc["intrinsics"]).reshape(-1, 10), torch.float64)
    model = d(sc["cam_model"], torch.int32) if cam_model else None
    off, cam, uv = d(view_off, torch.int32), d(view_cam, torch.int32), d(view_uv, torch.float64)
    n = view_off.size - 1
    xyz = torch.full((n, 3), float("nan"), dtype=torch.float64, device=dev)
    ok = torch.full((n,), 7, dtype=torch.uint8, device=dev) if want_ok else None
    p = lambda x: C.c_void_p(0 if x is None else x.data_ptr())
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.sba_triangulate_noalloc(sc["n_cams"], p(raxis), p(t), p(model), intr.shape[0], p(intr), n, int(view_off[-1]), p(off), p(cam), p(uv),
                                     p(xyz), p(ok), stream)
    assert rc == 0, lib.sba_create_error().decode()
    torch.cuda.synchronize()
    return xyz.cpu().numpy(), None if ok is None else ok.cpu().numpy().astype(bool)


def test_no_alloc_matches_the_existing_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    lib = _capi.load()
    assert lib.sba_triangulate_noalloc(0, None, None, None, 0, None, 0, 0, None, None, None, None, None, None) == 0   # set-up only
    big = problem(5, 64, 20000, 100000, two_models=True)
    small = problem(6, 24, 3000, 12000, two_models=False)          # fewer cameras / views: stale scratch of the first call stays behind it
    for sc in (big, small, big):
        voff, vcam, vuv = views_of(sc)
        x_ref, ok_ref = existing_entry_point(sc, voff, vcam, vuv)
        x, ok = no_alloc(torch, sc, voff, vcam, vuv)
        assert ok_ref.all() and ok.all()
        assert relerr(x, x_ref) < TOL
        if sc is small:          # sane geometry where the intrinsics are the true ones
            assert np.median(np.abs(x - sc["pts_true"]).max(axis=1)) < 0.5
    # cam_model = NULL means model 0 for every camera; ok_out may be NULL
    voff, vcam, vuv = views_of(small)
    x_ref, _ = existing_entry_point(small, voff, vcam, vuv)
    x, ok = no_alloc(torch, small, voff, vcam, vuv, cam_model=False, want_ok=False)
    assert ok is None and relerr(x, x_ref) < TOL


def test_no_alloc_limits_and_bad_indices():
    torch = pytest.importorskip("torch")
    lib = _capi.load()
    one = C.c_void_p(16)          # never dereferenced: the sizes are rejected first
    for args, what in (((4097, 10, 10), "4096"), ((10, 262145, 10), "262144"), ((10, 10, 1048577), "1048576")):
        rc = lib.sba_triangulate_noalloc(args[0], one, one, None, 1, one, args[1], args[2], one, one, one, one, None, None)
        assert rc == _capi.SBA_ERR_INVALID and what in lib.sba_create_error().decode()
    sc = problem(6, 24, 3000, 12000, two_models=False)
    voff, vcam, vuv = views_of(sc)
    x_ref, _ = existing_entry_point(sc, voff, vcam, vuv)
    vcam_bad = vcam.copy()
    hit = [5, 1700]
    vcam_bad[voff[hit[0]]] = 24; vcam_bad[voff[hit[1] + 1] - 1] = -3          # out of range: those two points fail, nothing else moves
    x, ok = no_alloc(torch, sc, voff, vcam_bad, vuv)
    assert not ok[hit].any() and not x[hit].any()
    keep = np.ones(sc["n_pts"], bool); keep[hit] = False
    assert ok[keep].all() and relerr(x[keep], x_ref[keep]) < TOL
