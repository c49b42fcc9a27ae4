"""The reference's project files (".steviapproj") through libsba_b200.so (sba_project_* in include/sba_b200.h).

Project::load / save (libs/datablocks/project.cpp:93-221) and the data-block -> problem logic of the SBA modules are C++
inside the library (csrc/sba_project.cpp); this module allocates the arrays and returns a scene dict that
`synthetic.load_into` registers on an `SBASolver`.
"""
import ctypes as C
import os

import numpy as np

from . import _capi

FIX_CAMERA_INTERNAL = 1   # FixedParameter::CameraInternal


def _check(rc, what):
    if rc != _capi.SBA_OK:
        raise ValueError(f"{what}: {_capi.load().sba_project_error().decode()}")


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Project:
    """A loaded project file: `scene()` gives the problem, `write_results()` + `save()` store a solution the way the
    reference's modules do (writeResults, then Project::save)."""

    def __init__(self, path):
        self._lib = _capi.load()
        self._h = C.c_void_p()
        _check(self._lib.sba_project_open(os.fsencode(path), C.byref(self._h)), "sba_project_open")
        v = [C.c_int32() for _ in range(7)]
        _check(self._lib.sba_project_sizes(self._h, *[C.byref(x) for x in v]), "sba_project_sizes")
        (self.n_models, self.n_images, self.n_landmarks, self.n_obs, self.n_priors, self.n_intrinsics_priors_max,
         self.n_images_without_camera) = [x.value for x in v]
        self.model_id = np.empty(self.n_models, np.int64)
        self.image_id = np.empty(self.n_images, np.int64)
        self.landmark_id = np.empty(self.n_landmarks, np.int64)
        _check(self._lib.sba_project_ids(self._h, _ptr(self.model_id), _ptr(self.image_id), _ptr(self.landmark_id)), "sba_project_ids")

    def close(self):
        if self._h:
            self._lib.sba_project_close(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def scene(self, fixed_flags=0):
        """Scene dict (keys of synthetic.make_scene + the free-intrinsics keys).  fixed_flags: FIX_CAMERA_INTERNAL holds every
        intrinsic block constant (ModularSBASolver::setFixedParametersFlag)."""
        nm, ni, nl, no, npr, nip = self.n_models, self.n_images, self.n_landmarks, self.n_obs, self.n_priors, self.n_intrinsics_priors_max
        intr = np.empty((nm, 10)); mask = np.empty(nm, np.uint8)
        raxis = np.empty((ni, 3)); t = np.empty((ni, 3)); cam_model = np.empty(ni, np.int32); cam_fixed = np.empty(ni, np.uint8)
        pts = np.empty((nl, 3)); pt_fixed = np.empty(nl, np.uint8)
        oc = np.empty(no, np.int32); op = np.empty(no, np.int32); uv = np.empty((no, 2)); info = np.empty((no, 4))
        pk = np.empty(npr, np.int32); pi = np.empty(npr, np.int32); pA = np.empty((npr, 9)); pb = np.empty((npr, 3))
        im = np.empty(nip, np.int32); ib = np.empty(nip, np.int32); iA = np.empty((nip, 9)); ibb = np.empty((nip, 3))
        n_ip = C.c_int32()
        _check(self._lib.sba_project_convert(self._h, fixed_flags, _ptr(intr), _ptr(mask), _ptr(raxis), _ptr(t), _ptr(cam_model),
                                             _ptr(cam_fixed), _ptr(pts), _ptr(pt_fixed), _ptr(oc), _ptr(op), _ptr(uv), _ptr(info),
                                             _ptr(pk), _ptr(pi), _ptr(pA), _ptr(pb), _ptr(im), _ptr(ib), _ptr(iA), _ptr(ibb),
                                             C.byref(n_ip)), "sba_project_convert")
        k = n_ip.value
        sc = {
            "n_cams": ni, "n_pts": nl, "n_obs": no,
            "intrinsics": intr, "cam_model": cam_model, "cam_fixed": cam_fixed, "cost_kind": np.zeros(ni, np.uint8),
            "pt_fixed": pt_fixed, "raxis": raxis, "t": t, "pts": pts,
            "obs_cam": oc, "obs_pt": op, "obs_uv": uv, "obs_info": info,
            "prior_kind": pk, "prior_index": pi, "prior_A": pA, "prior_b": pb,
        }
        if not np.all(mask == 0x1f):   # some intrinsic block is free: the reference's default
            sc["intr_fixed_mask"] = mask
            if k:
                sc["intr_prior_model"] = im[:k].copy(); sc["intr_prior_block"] = ib[:k].copy()
                sc["intr_prior_A"] = iA[:k].copy(); sc["intr_prior_b"] = ibb[:k].copy()
        return sc

    def write_results(self, raxis=None, t=None, pts=None, intrinsics=None):
        def arr(a):
            return None if a is None else np.ascontiguousarray(a, np.float64)
        raxis, t, pts, intrinsics = arr(raxis), arr(t), arr(pts), arr(intrinsics)
        _check(self._lib.sba_project_write_results(self._h, _ptr(raxis), _ptr(t), _ptr(pts), _ptr(intrinsics)), "sba_project_write_results")

    def save(self, path):
        _check(self._lib.sba_project_save(self._h, os.fsencode(path)), "sba_project_save")


def load_project(path, fixed_flags=0):
    """Scene dict of a project file (the handle is closed; use `Project` to write results back)."""
    with Project(path) as p:
        return p.scene(fixed_flags)
