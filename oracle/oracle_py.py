"""ctypes binding of oracle/liboracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm may import this.
The product (steviapp_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "liboracle.so")
MAX_HISTORY = 512


class OracleProblem(C.Structure):
    _fields_ = [
        ("n_cams", C.c_int32), ("n_pts", C.c_int32), ("n_obs", C.c_int32), ("n_models", C.c_int32), ("n_priors", C.c_int32),
        ("cam_r", C.c_void_p), ("cam_t", C.c_void_p), ("cam_model", C.c_void_p), ("cam_fixed", C.c_void_p),
        ("cam_cost_kind", C.c_void_p), ("pts", C.c_void_p), ("pt_fixed", C.c_void_p), ("intrinsics", C.c_void_p),
        ("obs_cam", C.c_void_p), ("obs_pt", C.c_void_p), ("obs_uv", C.c_void_p), ("obs_info", C.c_void_p),
        ("prior_kind", C.c_void_p), ("prior_index", C.c_void_p), ("prior_A", C.c_void_p), ("prior_b", C.c_void_p),
    ]


class OracleOptions(C.Structure):
    _fields_ = [
        ("max_iterations", C.c_int32),
        ("function_tolerance", C.c_double), ("parameter_tolerance", C.c_double), ("gradient_tolerance", C.c_double),
        ("initial_radius", C.c_double), ("max_radius", C.c_double), ("min_radius", C.c_double),
        ("min_lm_diagonal", C.c_double), ("max_lm_diagonal", C.c_double), ("min_relative_decrease", C.c_double),
        ("jacobi_scaling", C.c_int32), ("pcg_rel_tol", C.c_double), ("pcg_max_iters", C.c_int32),
        ("linear_solver", C.c_int32), ("num_threads", C.c_int32),
    ]


class OracleSummary(C.Structure):
    _fields_ = [
        ("status", C.c_int32), ("iterations", C.c_int32), ("successful_steps", C.c_int32),
        ("pcg_iterations_total", C.c_int32), ("initial_cost", C.c_double), ("final_cost", C.c_double),
        ("cost_history", C.c_double * MAX_HISTORY), ("radius_history", C.c_double * MAX_HISTORY),
        ("pcg_iters_history", C.c_int32 * MAX_HISTORY),
        ("seconds_total", C.c_double), ("seconds_evaluate", C.c_double), ("seconds_linear", C.c_double),
    ]


_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        _lib = C.CDLL(LIB)
        _lib.oracle_num_threads.restype = C.c_int
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Problem:
    """Holds numpy arrays (copied from a scene dict) and the C struct pointing into them."""

    def __init__(self, scene):
        g = lambda k, dt: None if scene.get(k) is None else np.ascontiguousarray(scene[k], dtype=dt).copy()
        self.cam_r = g("raxis", np.float64); self.cam_t = g("t", np.float64); self.pts = g("pts", np.float64)
        self.cam_model = g("cam_model", np.int32); self.cam_fixed = g("cam_fixed", np.uint8)
        self.cost_kind = g("cost_kind", np.uint8); self.pt_fixed = g("pt_fixed", np.uint8)
        self.intr = g("intrinsics", np.float64)
        self.obs_cam = g("obs_cam", np.int32); self.obs_pt = g("obs_pt", np.int32)
        self.obs_uv = g("obs_uv", np.float64); self.obs_info = g("obs_info", np.float64)
        self.prior_kind = g("prior_kind", np.int32); self.prior_index = g("prior_index", np.int32)
        self.prior_A = g("prior_A", np.float64); self.prior_b = g("prior_b", np.float64)
        self.n_cams = self.cam_r.size // 3; self.n_pts = self.pts.size // 3; self.n_obs = self.obs_cam.size
        n_pr = 0 if self.prior_kind is None else self.prior_kind.size
        self.c = OracleProblem(self.n_cams, self.n_pts, self.n_obs, self.intr.size // 10, n_pr,
                               _p(self.cam_r), _p(self.cam_t), _p(self.cam_model), _p(self.cam_fixed), _p(self.cost_kind),
                               _p(self.pts), _p(self.pt_fixed), _p(self.intr), _p(self.obs_cam), _p(self.obs_pt),
                               _p(self.obs_uv), _p(self.obs_info), _p(self.prior_kind), _p(self.prior_index),
                               _p(self.prior_A), _p(self.prior_b))

    def evaluate(self, threads=0, jac=True):
        lib = load()
        cost = C.c_double()
        r = np.empty((self.n_obs, 2)); jc = np.empty((self.n_obs, 2, 6)) if jac else None
        jp = np.empty((self.n_obs, 2, 3)) if jac else None
        failed = np.zeros(self.n_obs, dtype=np.uint8)
        rc = lib.oracle_evaluate(C.byref(self.c), threads, C.byref(cost), _p(r), _p(jc), _p(jp), _p(failed))
        return {"cost": cost.value, "ok": rc == 0, "residuals": r, "jac_cam": jc, "jac_pt": jp, "failed": failed}

    def intrinsics_jac(self):
        """d res / d (f, pp, k, t, B) per observation (n_obs, 2, 10): autodiff of the reference functor over the five intrinsic blocks."""
        lib = load()
        j = np.empty((self.n_obs, 2, 10))
        lib.oracle_evaluate_intrinsics_jac(C.byref(self.c), _p(j))
        return j

    def normal_blocks(self, threads=0):
        lib = load()
        gc = np.empty((self.n_cams, 6)); gp = np.empty((self.n_pts, 3))
        Hcc = np.empty((self.n_cams, 6, 6)); Hpp = np.empty((self.n_pts, 3, 3))
        rc = lib.oracle_normal_blocks(C.byref(self.c), threads, _p(gc), _p(gp), _p(Hcc), _p(Hpp))
        assert rc == 0
        return {"gc": gc, "gp": gp, "Hcc": Hcc, "Hpp": Hpp}

    def schur_matvec(self, lam_c, lam_p, x, threads=0):
        lib = load()
        lam_c = np.ascontiguousarray(lam_c, dtype=np.float64); lam_p = np.ascontiguousarray(lam_p, dtype=np.float64)
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty((self.n_cams, 6)); rhs = np.empty((self.n_cams, 6))
        rc = lib.oracle_schur_matvec(C.byref(self.c), threads, _p(lam_c), _p(lam_p), _p(x), _p(y), _p(rhs))
        assert rc == 0, rc
        return y, rhs

    def solve(self, **kw):
        lib = load()
        o = OracleOptions()
        lib.oracle_default_options(C.byref(o))
        for k, v in kw.items():
            if not hasattr(o, k):
                raise AttributeError(k)
            setattr(o, k, v)
        s = OracleSummary()
        lib.oracle_solve(C.byref(self.c), C.byref(o), C.byref(s))
        n = min(s.iterations, MAX_HISTORY - 1) + 1
        return {"status": s.status, "iterations": s.iterations, "successful_steps": s.successful_steps,
                "pcg_iterations_total": s.pcg_iterations_total, "initial_cost": s.initial_cost,
                "final_cost": s.final_cost, "cost_history": list(s.cost_history[:n]),
                "radius_history": list(s.radius_history[:n]), "pcg_iters_history": list(s.pcg_iters_history[:n]),
                "seconds_total": s.seconds_total, "seconds_evaluate": s.seconds_evaluate,
                "seconds_linear": s.seconds_linear}

    def triangulate(self, view_off, view_cam, view_uv, threads=0):
        lib = load()
        view_off = np.ascontiguousarray(view_off, dtype=np.int32); view_cam = np.ascontiguousarray(view_cam, dtype=np.int32)
        view_uv = np.ascontiguousarray(view_uv, dtype=np.float64)
        n = view_off.size - 1
        xyz = np.empty((n, 3)); ok = np.zeros(n, dtype=np.uint8)
        lib.oracle_triangulate(C.byref(self.c), n, _p(view_off), _p(view_cam), _p(view_uv), threads, _p(xyz), _p(ok))
        return xyz, ok.astype(bool)


def dir_from_uv(uv, intr10):
    lib = load()
    uv = np.ascontiguousarray(uv, dtype=np.float64); intr10 = np.ascontiguousarray(intr10, dtype=np.float64)
    d = np.empty(3)
    lib.oracle_dir_from_uv(_p(uv), _p(intr10), _p(d))
    return d


def uv_from_dir(xn, intr10):
    lib = load()
    xn = np.ascontiguousarray(xn, dtype=np.float64); intr10 = np.ascontiguousarray(intr10, dtype=np.float64)
    uv = np.empty(2)
    rc = lib.oracle_uv_from_dir(_p(xn), _p(intr10), _p(uv))
    assert rc == 0
    return uv


def num_threads():
    return load().oracle_num_threads()


def set_max_threads(n=0):
    """Caps the OpenMP team at n threads (0 = all online processors) whatever OMP_NUM_THREADS the launcher exported."""
    lib = load()
    lib.oracle_set_max_threads.restype = C.c_int
    return lib.oracle_set_max_threads(int(n))


# ---------------------------------------------------------------------------------------------------------------------
# extended problem class (sba_oracle_ext.cpp): free intrinsics, sensor offsets / lever arms, UV <-> UV factors, covariance
class OracleExtProblem(C.Structure):
    _fields_ = [
        ("base", OracleProblem),
        ("intr_rw", C.c_void_p), ("intr_fixed_mask", C.c_void_p),
        ("n_intr_priors", C.c_int32), ("ip_model", C.c_void_p), ("ip_block", C.c_void_p), ("ip_A", C.c_void_p), ("ip_b", C.c_void_p),
        ("cam_offset", C.c_void_p), ("cam_mount", C.c_void_p),
        ("n_mounts", C.c_int32), ("mount_b", C.c_void_p), ("mount_l", C.c_void_p), ("mount_fixed", C.c_void_p),
        ("n_pairs", C.c_int32), ("pair_cam", C.c_void_p), ("pair_uv", C.c_void_p), ("pair_info", C.c_void_p),
        ("loss_kind", C.c_int32), ("loss_a", C.c_double),
    ]


LOSS = {None: 0, "none": 0, "huber": 1, "cauchy": 2, "arctan": 3}


class ExtProblem(Problem):
    """Scene dict as for Problem plus the optional keys: intr_fixed_mask (n_models uint8; absent = intrinsics constant),
    intr_prior_model / _block / _A (n,9) / _b (n,3), cam_offset (n_cams,6), cam_mount (n_cams int32, -1 none), mount_b /
    mount_l (n_mounts,3), mount_fixed, pair_cam (n,2), pair_uv (n,4), pair_info (n,8), loss ("huber" | "cauchy" | "arctan"), loss_a."""

    def __init__(self, scene):
        super().__init__(scene)
        g = lambda k, dt: None if scene.get(k) is None else np.ascontiguousarray(scene[k], dtype=dt).copy()
        self.intr_mask = g("intr_fixed_mask", np.uint8)
        self.ip_model = g("intr_prior_model", np.int32); self.ip_block = g("intr_prior_block", np.int32)
        self.ip_A = g("intr_prior_A", np.float64); self.ip_b = g("intr_prior_b", np.float64)
        self.cam_offset = g("cam_offset", np.float64); self.cam_mount = g("cam_mount", np.int32)
        self.mount_b = g("mount_b", np.float64); self.mount_l = g("mount_l", np.float64); self.mount_fixed = g("mount_fixed", np.uint8)
        self.pair_cam = g("pair_cam", np.int32); self.pair_uv = g("pair_uv", np.float64); self.pair_info = g("pair_info", np.float64)
        n_ip = 0 if self.ip_model is None else self.ip_model.size
        n_mt = 0 if self.mount_b is None else self.mount_b.size // 3
        n_pr = 0 if self.pair_cam is None else self.pair_cam.size // 2
        self.n_mounts, self.n_pairs = n_mt, n_pr
        self.e = OracleExtProblem(self.c, _p(self.intr) if self.intr_mask is not None else None, _p(self.intr_mask),
                                  n_ip, _p(self.ip_model), _p(self.ip_block), _p(self.ip_A), _p(self.ip_b),
                                  _p(self.cam_offset), _p(self.cam_mount), n_mt, _p(self.mount_b), _p(self.mount_l), _p(self.mount_fixed),
                                  n_pr, _p(self.pair_cam), _p(self.pair_uv), _p(self.pair_info),
                                  LOSS[scene.get("loss")], float(scene.get("loss_a", 1.0)))

    def layout(self):
        lib = load()
        n = C.c_int32()
        mb = np.zeros(max(self.intr.size // 10, 1), dtype=np.int32); tb = np.zeros(max(self.n_mounts, 1), dtype=np.int32)
        lib.oracle_ext_layout(C.byref(self.e), C.byref(n), _p(mb), _p(tb), None)
        live = np.zeros(6 * n.value, dtype=np.uint8)
        lib.oracle_ext_layout(C.byref(self.e), C.byref(n), _p(mb), _p(tb), _p(live))
        return {"n_ext": n.value, "model_block": mb[:self.intr.size // 10], "mount_block": tb[:self.n_mounts], "live": live.astype(bool)}

    def evaluate(self, threads=0, jac=True):
        lib = load()
        cost = C.c_double()
        n, k = self.n_obs, self.n_pairs
        r = np.empty((n, 2)); jc = np.empty((n, 2, 6)); jp = np.empty((n, 2, 3)); ji = np.empty((n, 2, 10)); jl = np.empty((n, 2, 6))
        pr = np.empty((k, 4)); pj = np.empty((k, 4, 32))
        rc = lib.oracle_ext_evaluate(C.byref(self.e), C.byref(cost), _p(r), _p(jc), _p(jp), _p(ji), _p(jl), _p(pr), _p(pj))
        return {"cost": cost.value, "ok": rc == 0, "residuals": r, "jac_cam": jc, "jac_pt": jp, "jac_intr": ji, "jac_lever": jl,
                "pair_residuals": pr, "pair_jac": pj}

    def _options(self, kw):
        lib = load()
        o = OracleOptions()
        lib.oracle_default_options(C.byref(o))
        for k, v in kw.items():
            if not hasattr(o, k):
                raise AttributeError(k)
            setattr(o, k, v)
        return o

    def reduced(self, radius=0.0, **kw):
        lib = load()
        n = 6 * self.layout()["n_ext"]
        S = np.empty((n, n)); rhs = np.empty(n)
        o = self._options(kw)
        rc = lib.oracle_ext_reduced(C.byref(self.e), C.byref(o), C.c_double(radius), _p(S), _p(rhs))
        assert rc == 0, rc
        return S, rhs

    def point_covariance(self):
        lib = load()
        cov = np.empty((self.n_pts, 3, 3))
        rc = lib.oracle_ext_point_covariance(C.byref(self.e), _p(cov))
        assert rc == 0, rc
        return cov

    def solve(self, **kw):
        lib = load()
        o = self._options(kw)
        s = OracleSummary()
        lib.oracle_ext_solve(C.byref(self.e), C.byref(o), C.byref(s))
        n = min(s.iterations, MAX_HISTORY - 1) + 1
        return {"status": s.status, "iterations": s.iterations, "successful_steps": s.successful_steps,
                "pcg_iterations_total": s.pcg_iterations_total, "initial_cost": s.initial_cost, "final_cost": s.final_cost,
                "cost_history": list(s.cost_history[:n]), "radius_history": list(s.radius_history[:n]),
                "pcg_iters_history": list(s.pcg_iters_history[:n]), "seconds_total": s.seconds_total}
