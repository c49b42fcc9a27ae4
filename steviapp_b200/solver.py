"""Host-side mirror of the reference's solver interface over the C ABI.

`SBASolver` follows ModularSBASolver's life cycle
(libs/sparsesolver/modularsbasolver.{h,cpp}): parameter blocks are registered
by reference and updated IN PLACE by the solve (numpy arrays here, raw
`double*` in the reference), observations are added through the bulk form of
ProjectorModule::addProjectionCostFunction (modularsbasolver.h:169-177),
`init()` / `opt_step()` / `cleanup()` keep the SteppedProcess protocol
(libs/processing/steppedprocess.cpp:58-115, sparsesolverbase.cpp:71-88): the
first `opt_step()` runs the whole LM, later calls return True immediately
(modularsbasolver.cpp:886-888).

All arithmetic happens in libsba_b200.so on the GPU; nothing here computes.
"""
import ctypes as C

import numpy as np

from . import _capi


class SBAError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"sba_b200 error {code}: {msg}")
        self.code = code


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        a = a.reshape(shape)
    return a


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def default_options(**kw):
    lib = _capi.load()
    o = _capi.SbaOptions()
    lib.sba_default_options(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise AttributeError(f"sba_options has no field {k}")
        setattr(o, k, v)
    return o


def summary_to_dict(s):
    n = min(s.iterations, _capi.SBA_MAX_HISTORY - 1) + 1
    return {
        "status": s.status, "iterations": s.iterations, "successful_steps": s.successful_steps,
        "pcg_iterations_total": s.pcg_iterations_total,
        "initial_cost": s.initial_cost, "final_cost": s.final_cost,
        "cost_history": list(s.cost_history[:n]), "radius_history": list(s.radius_history[:n]),
        "pcg_iters_history": list(s.pcg_iters_history[:n]),
        "ms_total": s.ms_total, "ms_linearize": s.ms_linearize, "ms_schur": s.ms_schur, "ms_pcg": s.ms_pcg,
        "ms_update": s.ms_update, "ms_h2d": s.ms_h2d, "ms_d2h": s.ms_d2h, "kernel_launches": s.kernel_launches,
    }


class SBASolver:
    """One handle = one rank's shard of a bundle-adjustment problem, resident on one GPU."""

    def __init__(self, options=None, **kw):
        self._lib = _capi.load()
        self._opt = options if options is not None else default_options(**kw)
        h = C.c_void_p()
        rc = self._lib.sba_create(C.byref(self._opt), C.byref(h))
        if rc != 0:
            raise SBAError(rc, self._lib.sba_create_error().decode())
        self._h = h
        self._keep = {}          # arrays whose memory the library points into / that must outlive calls
        self._not_first_step = False
        self._cb = None
        self.last_summary = None

    # -- plumbing ---------------------------------------------------------
    def _check(self, rc, allow=()):
        if rc != 0 and rc not in allow:
            raise SBAError(rc, self._lib.sba_last_error(self._h).decode())
        return rc

    def close(self):
        if getattr(self, "_h", None):
            self._lib.sba_destroy(self._h)
            self._h = None

    cleanup = close   # ModularSBASolver::cleanup

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- problem construction (init() in the reference) ---------------------
    def set_options(self, **kw):
        for k, v in kw.items():
            if not hasattr(self._opt, k):
                raise AttributeError(f"sba_options has no field {k}")
            setattr(self._opt, k, v)
        self._check(self._lib.sba_set_options(self._h, C.byref(self._opt)))

    def set_intrinsics(self, intr):
        intr = _f64(intr).reshape(-1, 10)
        self._keep["intr"] = intr
        self._check(self._lib.sba_set_intrinsics(self._h, intr.shape[0], _ptr(intr)))

    def set_intrinsics_blocks(self, intr, fixed_mask=None):
        """Camera models whose blocks f, pp, k, t, B may be FREE (the reference's default).  intr: (n_models, 10) float64
        C-contiguous, UPDATED IN PLACE by solve(); fixed_mask: per model, bit k set -> block k constant (None: all constant)."""
        if not (isinstance(intr, np.ndarray) and intr.dtype == np.float64 and intr.flags.c_contiguous and intr.flags.writeable):
            raise TypeError("intr must be a writeable C-contiguous float64 ndarray (it is updated in place)")
        n = intr.size // 10
        fm = None if fixed_mask is None else np.ascontiguousarray(fixed_mask, dtype=np.uint8)
        if fm is not None and fm.size != n:
            raise ValueError("fixed_mask must have one entry per camera model")
        self._keep.update(intr=intr, intr_mask=fm)
        self._check(self._lib.sba_set_intrinsics_blocks(self._h, n, _ptr(intr), _ptr(fm)))

    def add_intrinsics_priors(self, model, block, A, b):
        """FixedSizeNormalPrior<n,n> on intrinsic blocks: A (n, 9) with dim x dim row-major leading entries, b (n, 3)."""
        model = np.ascontiguousarray(model, dtype=np.int32); block = np.ascontiguousarray(block, dtype=np.int32)
        A = _f64(A); b = _f64(b)
        if block.size != model.size or A.size != 9 * model.size or b.size != 3 * model.size:
            raise ValueError("intrinsics priors: model/block/A/b sizes disagree")
        self._check(self._lib.sba_add_intrinsics_priors(self._h, model.size, _ptr(model), _ptr(block), _ptr(A), _ptr(b)))

    def set_sensor_offsets(self, offset6):
        """Constant sensor -> body transform per camera, (n_cams, 6) = (r_off, t_off); None = identity."""
        off = None if offset6 is None else _f64(offset6)
        self._check(self._lib.sba_set_sensor_offsets(self._h, _ptr(off)))

    def set_lever_arms(self, boresight, lever, cam_mount, fixed=None):
        """Free boresight / lever-arm blocks: boresight, lever (n_mounts, 3) float64, UPDATED IN PLACE; cam_mount (n_cams) int32, -1 = none."""
        for name, a in (("boresight", boresight), ("lever", lever)):
            if not (isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags.c_contiguous and a.flags.writeable):
                raise TypeError(f"{name} must be a writeable C-contiguous float64 ndarray (it is updated in place)")
        n = boresight.size // 3
        cm = np.ascontiguousarray(cam_mount, dtype=np.int32)
        fx = None if fixed is None else np.ascontiguousarray(fixed, dtype=np.uint8)
        self._keep.update(mount_b=boresight, mount_l=lever, cam_mount=cm, mount_fixed=fx)
        self._check(self._lib.sba_set_lever_arms(self._h, n, _ptr(boresight), _ptr(lever), _ptr(fx), _ptr(cm)))

    def add_uv_pairs(self, cam1, cam2, uv1, uv2, info1=None, info2=None):
        """UV <-> UV cross-projection factors (addCrossProjectionCostFunction): uv1, uv2 (n, 2), info (n, 2, 2) or None."""
        cam1 = np.ascontiguousarray(cam1, dtype=np.int32); cam2 = np.ascontiguousarray(cam2, dtype=np.int32)
        uv1 = _f64(uv1); uv2 = _f64(uv2)
        i1 = None if info1 is None else _f64(info1); i2 = None if info2 is None else _f64(info2)
        if cam2.size != cam1.size or uv1.size != 2 * cam1.size or uv2.size != 2 * cam1.size:
            raise ValueError("uv pairs: sizes disagree")
        self._n_pairs = getattr(self, "_n_pairs", 0) + cam1.size
        self._check(self._lib.sba_add_uv_pairs(self._h, cam1.size, _ptr(cam1), _ptr(cam2), _ptr(uv1), _ptr(uv2), _ptr(i1), _ptr(i2)))

    def set_pair_loss(self, kind, a=1.0):
        kinds = {None: 0, "none": 0, "huber": 1, "cauchy": 2, "arctan": 3}
        self._check(self._lib.sba_set_pair_loss(self._h, kinds.get(kind, kind), float(a)))

    def evaluate_pairs(self):
        n = getattr(self, "_n_pairs", 0)
        r = np.empty((n, 4)); j = np.empty((n, 4, 12))
        self._check(self._lib.sba_evaluate_pairs(self._h, _ptr(r), _ptr(j)))
        return r, j

    def point_covariances(self):
        """std_step: (n_pts, 3, 3) marginal covariances of the point blocks at the current (device-resident) parameters."""
        self.finalize()
        cov = np.empty((self.sizes()["n_pts"], 3, 3))
        self._check(self._lib.sba_point_covariances(self._h, _ptr(cov)))
        return cov

    def shared_info(self):
        """Layout of the reduced camera system: extended cameras, groups, first pseudo-camera of every model / mounting (-1: constant)."""
        n_ext, n_g = C.c_int32(), C.c_int32()
        nm = self._keep["intr"].size // 10 if "intr" in self._keep else 0
        nt = self._keep["mount_b"].size // 3 if "mount_b" in self._keep else 0
        mb = np.full(max(nm, 1), -1, dtype=np.int32); tb = np.full(max(nt, 1), -1, dtype=np.int32)
        self._check(self._lib.sba_shared_info(self._h, C.byref(n_ext), C.byref(n_g), _ptr(mb), _ptr(tb)))
        return {"n_ext": n_ext.value, "n_groups": n_g.value, "model_block": mb[:nm], "mount_block": tb[:nt]}

    def set_poses(self, raxis, t, cam_model=None, fixed=None, cost_kind=None):
        """raxis, t: (n,3) float64 C-contiguous arrays, sensor->world; UPDATED IN PLACE by solve()."""
        for name, a in (("raxis", raxis), ("t", t)):
            if not (isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags.c_contiguous and a.flags.writeable):
                raise TypeError(f"{name} must be a writeable C-contiguous float64 ndarray (it is updated in place)")
        n = raxis.size // 3
        if t.size != 3 * n:
            raise ValueError("raxis and t must have the same number of poses")
        cm = None if cam_model is None else np.ascontiguousarray(cam_model, dtype=np.int32)
        fx = None if fixed is None else np.ascontiguousarray(fixed, dtype=np.uint8)
        ck = None if cost_kind is None else np.ascontiguousarray(cost_kind, dtype=np.uint8)
        for a in (cm, fx, ck):
            if a is not None and a.size != n:
                raise ValueError("per-pose arrays must have n entries")
        self._keep.update(raxis=raxis, t=t, cam_model=cm, cam_fixed=fx, cam_kind=ck)
        self._check(self._lib.sba_set_poses(self._h, _ptr(raxis), _ptr(t), n, _ptr(cm), _ptr(fx), _ptr(ck)))

    def set_points(self, xyz, fixed=None):
        """xyz: (n,3) float64 C-contiguous; UPDATED IN PLACE by solve()."""
        if not (isinstance(xyz, np.ndarray) and xyz.dtype == np.float64 and xyz.flags.c_contiguous and xyz.flags.writeable):
            raise TypeError("xyz must be a writeable C-contiguous float64 ndarray (it is updated in place)")
        n = xyz.size // 3
        fx = None if fixed is None else np.ascontiguousarray(fixed, dtype=np.uint8)
        if fx is not None and fx.size != n:
            raise ValueError("fixed must have n entries")
        self._keep.update(xyz=xyz, pt_fixed=fx)
        self._check(self._lib.sba_set_points(self._h, _ptr(xyz), n, _ptr(fx)))

    def add_observations(self, cam_idx, pt_idx, uv, info=None):
        """Bulk ProjectorModule::addProjectionCostFunction: uv (n,2), info (n,2,2) or None (identity)."""
        cam_idx = np.ascontiguousarray(cam_idx, dtype=np.int32)
        pt_idx = np.ascontiguousarray(pt_idx, dtype=np.int32)
        uv = _f64(uv)
        n = cam_idx.size
        if pt_idx.size != n or uv.size != 2 * n:
            raise ValueError("cam_idx, pt_idx and uv must describe the same number of observations")
        inf = None
        if info is not None:
            inf = _f64(info)
            if inf.size != 4 * n:
                raise ValueError("info must hold one 2x2 matrix per observation")
        self._check(self._lib.sba_add_observations(self._h, n, _ptr(cam_idx), _ptr(pt_idx), _ptr(uv), _ptr(inf)))

    def add_projection_cost_function(self, pt_index, cam_index, uv, stiffness=None):
        """One-at-a-time form (same argument meaning as addProjectionCostFunction; indices instead of pointers)."""
        self.add_observations([cam_index], [pt_index], np.asarray(uv, dtype=np.float64).reshape(1, 2),
                              None if stiffness is None else np.asarray(stiffness, dtype=np.float64).reshape(1, 4))
        return True

    def add_priors(self, kind, index, A, b):
        kind = np.ascontiguousarray(kind, dtype=np.int32)
        index = np.ascontiguousarray(index, dtype=np.int32)
        A = _f64(A)
        b = _f64(b)
        n = kind.size
        if index.size != n or A.size != 9 * n or b.size != 3 * n:
            raise ValueError("priors: kind/index/A/b sizes disagree")
        self._check(self._lib.sba_add_priors(self._h, n, _ptr(kind), _ptr(index), _ptr(A), _ptr(b)))

    def finalize(self):
        self._check(self._lib.sba_finalize(self._h))

    def init(self):
        """ModularSBASolver::init: returns True on success (False + message in the reference's failed() path)."""
        self.finalize()
        return True

    def sizes(self):
        v = [C.c_int32() for _ in range(5)]
        self._check(self._lib.sba_problem_sizes(self._h, *[C.byref(x) for x in v]))
        return dict(zip(("n_cams", "n_pts", "n_obs", "n_slots", "n_tiles"), [x.value for x in v]))

    def intrinsics_jacobian(self):
        """d res / d (f, pp[2], k[3], t[2], B[2]) per observation, (n_obs, 2, 10), caller order, at the current parameters."""
        self.finalize()
        j = np.empty((self.sizes()["n_obs"], 2, 10))
        self._check(self._lib.sba_evaluate_intrinsics_jacobian(self._h, _ptr(j)))
        return j

    def evaluate_offset(self, off12):
        """Residuals / Jacobians with a constant sensor offset per camera (off12: (n_cams, 12) = Roff^T row-major + toff);
        returns dict(residuals (n_obs, 2), jac_cam (n_obs, 2, 6), jac_pt (n_obs, 2, 3), n_failed), caller order."""
        self.finalize()
        n = self.sizes()["n_obs"]
        off12 = np.ascontiguousarray(off12, dtype=np.float64)
        assert off12.shape == (self.sizes()["n_cams"], 12)
        r = np.empty((n, 2)); jc = np.empty((n, 2, 6)); jp = np.empty((n, 2, 3)); nf = C.c_int32()
        self._check(self._lib.sba_evaluate_offset(self._h, _ptr(off12), _ptr(r), _ptr(jc), _ptr(jp), C.byref(nf)))
        return {"residuals": r, "jac_cam": jc, "jac_pt": jp, "n_failed": nf.value}

    def explicit_info(self):
        """Static sizes of the explicit reduced camera system (enabled = the PCG applies S from the tiles)."""
        v = [C.c_int32() for _ in range(4)]
        self._check(self._lib.sba_explicit_info(self._h, *[C.byref(x) for x in v]))
        return dict(zip(("enabled", "n_tiles", "n_blocks", "n_contributions"), [x.value for x in v]))

    def explicit_shard(self):
        """This rank's slice of the explicit system's tiles and whether the PCG all-reduce runs over peer mailboxes."""
        v = [C.c_int32() for _ in range(4)]
        self._check(self._lib.sba_explicit_shard(self._h, *[C.byref(x) for x in v]))
        return dict(zip(("first_tile", "n_own", "peer_exchange", "persistent_pcg"), [x.value for x in v]))

    def explicit_sparse_info(self):
        """Block-sparse copy of the tiles read by the persistent PCG kernel (panels of non-zero 6x6 blocks, chunks = bulk copies)."""
        v = [C.c_int32(), C.c_int32(), C.c_int32(), C.c_int64(), C.c_int64()]
        self._check(self._lib.sba_explicit_sparse_info(self._h, *[C.byref(x) for x in v]))
        return dict(zip(("enabled", "n_panels", "n_chunks", "n_blocks", "stream_bytes"), [x.value for x in v]))

    def explicit_dense(self):
        """The reduced camera system of the last LM iteration as a dense symmetric matrix (6 n_cams)^2 and its rhs
        (inspection hook over sba_explicit_tiles; small problems / studies only)."""
        info = self.explicit_info()
        if not info["enabled"]:
            raise SBAError(-1, "explicit reduced camera system is off")
        nt, nc = info["n_tiles"], max(self.sizes()["n_cams"], self.shared_info()["n_ext"])
        tiles = np.empty((nt, 384, 384)); ti = np.empty(nt, dtype=np.int32); tj = np.empty(nt, dtype=np.int32)
        rhs = np.empty(6 * nc)
        self._check(self._lib.sba_explicit_tiles(self._h, _ptr(tiles), _ptr(ti), _ptr(tj), _ptr(rhs)))
        NT = (nc + 63) // 64
        U = np.zeros((NT * 384, NT * 384))
        for t in range(nt):
            U[384 * ti[t]:384 * ti[t] + 384, 384 * tj[t]:384 * tj[t] + 384] = tiles[t]
        S = U + U.T          # diagonal tiles store S_ii / 2 and nothing below: the sum restores the symmetric matrix
        return S[:6 * nc, :6 * nc], rhs

    # -- evaluation hooks -------------------------------------------------
    def upload_parameters(self):
        self._check(self._lib.sba_upload_parameters(self._h))

    def download_parameters(self):
        self._check(self._lib.sba_download_parameters(self._h))

    def evaluate(self, residuals=True, jacobians=True):
        """ceres::Problem::Evaluate equivalent; returns dict(cost, ok, residuals, jac_cam, jac_pt) in caller order."""
        self.finalize()
        n = self.sizes()["n_obs"]
        cost = C.c_double()
        r = np.empty((n, 2)) if residuals else None
        jc = np.empty((n, 2, 6)) if jacobians else None
        jp = np.empty((n, 2, 3)) if jacobians else None
        rc = self._check(self._lib.sba_evaluate(self._h, C.byref(cost), _ptr(r), _ptr(jc), _ptr(jp)), allow=(1,))
        return {"cost": cost.value, "ok": rc == 0, "residuals": r, "jac_cam": jc, "jac_pt": jp}

    def normal_blocks(self):
        self.finalize()
        s = self.sizes()
        gc = np.empty((s["n_cams"], 6)); gp = np.empty((s["n_pts"], 3))
        Hcc = np.empty((s["n_cams"], 6, 6)); Hpp = np.empty((s["n_pts"], 3, 3))
        self._check(self._lib.sba_normal_blocks(self._h, _ptr(gc), _ptr(gp), _ptr(Hcc), _ptr(Hpp)))
        return {"gc": gc, "gp": gp, "Hcc": Hcc, "Hpp": Hpp}

    def schur_matvec(self, lam_c, lam_p, x):
        self.finalize()
        s = self.sizes()
        lam_c = _f64(lam_c); lam_p = _f64(lam_p); x = _f64(x)
        y = np.empty((s["n_cams"], 6)); rhs = np.empty((s["n_cams"], 6))
        self._check(self._lib.sba_schur_matvec(self._h, _ptr(lam_c), _ptr(lam_p), _ptr(x), _ptr(y), _ptr(rhs)))
        return y, rhs

    # -- solve --------------------------------------------------------------
    def set_iteration_callback(self, fn):
        """fn(info_dict) -> truthy to abort; replaces ModularSBASolverIterationCallback."""
        if fn is None:
            self._cb = None
            self._check(self._lib.sba_set_iteration_callback(self._h, _capi.ITERATION_CALLBACK(), None))
            return

        def tramp(info_p, _user):
            i = info_p.contents
            return 1 if fn({f[0]: getattr(i, f[0]) for f in _capi.SbaIterationInfo._fields_}) else 0

        self._cb = _capi.ITERATION_CALLBACK(tramp)
        self._check(self._lib.sba_set_iteration_callback(self._h, self._cb, None))

    def solve(self, max_iterations=None, copy_params=True, **opts):
        if max_iterations is not None:
            opts["max_iterations"] = int(max_iterations)
        if opts:
            self.set_options(**opts)
        s = _capi.SbaSummary()
        self._check(self._lib.sba_solve(self._h, 1 if copy_params else 0, C.byref(s)))
        self.last_summary = summary_to_dict(s)
        return self.last_summary

    def opt_step(self):
        """ModularSBASolver::opt_step: the first call runs the whole LM, later calls are no-ops returning True."""
        if self._not_first_step:
            return True
        try:
            self.solve()
        except SBAError:
            return False
        self._not_first_step = True
        return True

    # -- measurement / adjacent ops -------------------------------------
    def time_stage(self, stage, reps=10):
        ms = C.c_double()
        self._check(self._lib.sba_time_stage(self._h, int(stage), int(reps), C.byref(ms)))
        return ms.value

    def triangulate(self, view_off, view_cam, view_uv):
        view_off = np.ascontiguousarray(view_off, dtype=np.int32)
        view_cam = np.ascontiguousarray(view_cam, dtype=np.int32)
        view_uv = _f64(view_uv)
        n = view_off.size - 1
        xyz = np.empty((n, 3)); ok = np.empty(n, dtype=np.uint8)
        ms = C.c_double()
        self._check(self._lib.sba_triangulate(self._h, n, _ptr(view_off), _ptr(view_cam), _ptr(view_uv), _ptr(xyz),
                                               _ptr(ok), C.byref(ms)))
        return xyz, ok.astype(bool), ms.value

    def comm_init(self, rank, nranks, unique_id):
        uid = np.frombuffer(bytes(unique_id), dtype=np.uint8).copy()
        if uid.size != 128:
            raise ValueError("unique_id must be 128 bytes")
        self._check(self._lib.sba_comm_init(self._h, int(rank), int(nranks), _ptr(uid)))


def comm_unique_id():
    lib = _capi.load()
    buf = np.zeros(128, dtype=np.uint8)
    rc = lib.sba_comm_unique_id(_ptr(buf))
    if rc != 0:
        raise SBAError(rc, lib.sba_create_error().decode())
    return buf.tobytes()


def partition_points(obs_per_point, nranks):
    lib = _capi.load()
    opp = np.ascontiguousarray(obs_per_point, dtype=np.int32)
    begin = np.zeros(nranks + 1, dtype=np.int32)
    rc = lib.sba_partition_points(opp.size, _ptr(opp), int(nranks), _ptr(begin))
    if rc != 0:
        raise SBAError(rc, "sba_partition_points: invalid arguments")
    return begin
