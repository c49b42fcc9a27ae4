#!/usr/bin/env python
"""bench.py -- LM iterations/s (+ residual/Jacobian Mobs/s, HBM GB/s vs peak) on the synthetic
Venice-shape bundle adjustment (1778 cameras / 993 923 points / 5 001 946 observations).

  python bench.py --gpus N --steps K --warmup W            B200 engine (libsba_b200.so)
  python bench.py --impl reference --gpus N --steps K ...  the reference's CPU path (oracle port), host cores

A "step" is ONE Levenberg-Marquardt iteration: damping + point-block inversion + reduced rhs +
SCHUR_JACOBI preconditioner + explicit reduced camera system + PCG (one persistent kernel) + back-substitution + trial cost
(+ re-linearisation when the step is accepted).  Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "lm_iterations_per_s"
UNIT = "LM iters/s"
FIXED = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
LM_CHUNK = 6             # LM iterations per uninterrupted solve of the timed region (restart from the initial values in between)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--shape", default="venice")
    ap.add_argument("--pcg-tol", type=float, default=1e-10)
    ap.add_argument("--pcg-max-iters", type=int, default=500)
    ap.add_argument("--pcg-check-every", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--free-intrinsics", action="store_true", help="optimise the intrinsic blocks f, pp, k, t of the camera model too (the reference's default; SURVEY 8f row 1)")
    ap.add_argument("--mixed", action="store_true", help="cameras alternate plane / spherical cost (BASELINE config 4)")
    ap.add_argument("--dlt-points", type=float, default=0.0, help="also time the batched DLT triangulation of this many million points (BASELINE config 4: 10)")
    ap.add_argument("--eta", type=float, default=0.1, help="forcing term of the labelled inexact-Newton second figure (0: skip it)")
    ap.add_argument("--linear-solver", type=int, default=0, help="0 SCHUR_JACOBI PCG (headline), 1 dense Cholesky of the explicit reduced system")
    ap.add_argument("--no-single-thread", action="store_true", help="reference arm: skip the 1-thread LM iteration (the reference's real configuration)")
    return ap.parse_args()


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.stop = threading.Event()
        self.th = None

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([x.strip() for x in out.strip().split(",")])
            except Exception:
                pass
            self.stop.wait(0.01)

    def __enter__(self):
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=6)

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------- CPU legs (oracle)
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_leg(shape, steps, warmup, pcg_tol, pcg_max, threads=0, single_thread_probe=False, scene=None, budget_s=150.0):
    """Times the CPU restatement of the reference (oracle/) on the FULL workload -- the same scene, the same chunks of at
    most LM_CHUNK LM iterations restarting from the initial values as the B200 arm -- with an explicit thread count
    (all host cores by default, whatever OMP_NUM_THREADS a launcher exported).  value = LM iterations / wall second."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py
    from steviapp_b200 import synthetic
    nthreads = host_cores() if threads <= 0 else threads
    nthreads = oracle_py.set_max_threads(nthreads)
    sc = synthetic.make_scene(shape) if scene is None else scene
    kw = dict(num_threads=nthreads, pcg_rel_tol=pcg_tol, pcg_max_iters=pcg_max, **FIXED)
    if warmup > 0:
        oracle_py.Problem(sc).solve(max_iterations=warmup, **kw)
    dt, done, pcg, first = 0.0, 0, 0, None
    asked = steps
    while done < steps:
        n = min(LM_CHUNK, steps - done)
        if done and dt + dt / done * n > budget_s:   # bounded sample: stop at a chunk boundary once the budget would be exceeded
            steps = done
            break
        p = oracle_py.Problem(sc)
        t0 = time.perf_counter()
        s = p.solve(max_iterations=n, **kw)
        dt += time.perf_counter() - t0
        assert s["iterations"] == n, s
        done += n
        pcg += s["pcg_iterations_total"]
        if first is None:
            first = s
    t1 = time.perf_counter()
    oracle_py.Problem(sc).evaluate(threads=nthreads)
    t_eval = time.perf_counter() - t1
    out = {
        "value": steps / dt, "unit": UNIT, "cores": nthreads, "kind": "port",
        "sample": (f"the full {shape}-shape scene ({sc['n_cams']} cameras / {sc['n_pts']} points / {sc['n_obs']} observations), "
                   f"{steps} LM iterations in solves of at most {LM_CHUNK} (each from the initial values) in {dt:.2f} s wall, "
                   f"{pcg} PCG iterations, {nthreads} OpenMP threads; {warmup} untimed warm-up iteration(s)"
                   + ("" if steps == asked else f"; {asked} steps were asked for, the {budget_s:.0f} s CPU budget ended the sample at {steps}")),
        "seconds": dt, "steps": steps, "pcg_iterations": pcg,
        "cost_history_first_solve": first["cost_history"], "pcg_iters_first_solve": first["pcg_iters_history"],
        "residual_jacobian_mobs_per_s": sc["n_obs"] / t_eval / 1e6,
        "note": "CPU restatement of the reference (not Ceres): the reference itself needs Eigen/Ceres/Qt/LibStevi, absent here",
    }
    if single_thread_probe and dt / steps * nthreads * 0.7 > 150.0:
        out["single_thread"] = {"skipped": f"one 1-thread LM iteration would take > 150 s on this host ({dt / steps:.1f} s with {nthreads} threads)"}
    elif single_thread_probe:
        # the reference's real configuration: ceres::Solver::Options::num_threads is never set (modularsbasolver.cpp:897-924)
        oracle_py.set_max_threads(1)
        p = oracle_py.Problem(sc)
        t0 = time.perf_counter()
        s1 = p.solve(max_iterations=1, **dict(kw, num_threads=1))
        d1 = time.perf_counter() - t0
        oracle_py.set_max_threads(nthreads)
        out["single_thread"] = {"value": s1["iterations"] / d1, "unit": UNIT, "cores": 1,
                                "sample": f"1 LM iteration ({s1['pcg_iterations_total']} PCG iterations) of the same scene in {d1:.2f} s"}
    return out


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb = cpu_leg(a.shape, max(a.steps, 1), min(a.warmup, 1), a.pcg_tol, a.pcg_max_iters, single_thread_probe=not a.no_single_thread)
    v = cb["value"]
    cfg = bench_config(a)
    cfg["linear_solver"] = "Schur complement + SCHUR_JACOBI PCG on the implicit operator (CPU oracle, OpenMP)"
    cfg["parallelism"] = f"{cb['cores']} host threads (rank 0 only)"
    cfg["l2"] = "n/a (CPU arm)"
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": min(a.warmup, 1), "ms_per_step": 1e3 * cb["seconds"] / cb["steps"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": cfg, "cpu_baseline": cb, "final_cost": cb["cost_history_first_solve"][-1],
        "final_cost_after_lm_iterations": len(cb["cost_history_first_solve"]) - 1,
        "cost_history_first_solve": cb["cost_history_first_solve"],
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def bench_config(a, explicit=True):
    from steviapp_b200 import synthetic
    nc, npt, no = synthetic.SHAPES[a.shape]
    return {"workload": f"synthetic {a.shape}-shape BA: {nc} cameras / {npt} points / {no} observations, "
                        + ("cameras alternate plane / spherical cost, " if getattr(a, "mixed", False) else "plane cost, ")
                        + ("intrinsic blocks f, pp, k, t FREE (one shared model: two pseudo-cameras in the reduced system), " if getattr(a, "free_intrinsics", False) else "intrinsics constant, ")
                        + "first camera fixed, 4 GCP priors",
            "pcg_rel_tol": a.pcg_tol, "pcg_max_iters": a.pcg_max_iters, "jacobi_scaling": 1,
            "linear_solver": ("Schur complement + dense FP64 Cholesky of the explicit S on the device (sba_options.linear_solver = 1, the reference's SPARSE_SCHUR-style direct solve; hand-written blocked factorisation, csrc/sba_kernels_dense.cu)") if getattr(a, "linear_solver", 0) == 1 else ("Schur complement + SCHUR_JACOBI PCG; S explicit (dense symmetric 384x384 tiles, rebuilt every LM iteration; when a rank holds >= 64 MB of them the PCG streams them as block-sparse panels: only the 6x6 blocks of camera pairs that share a point); "
                              "whole PCG solve = one persistent cooperative kernel (TMA bulk-copy ring); multi-GPU: partial tiles reduce-scattered, "
                              "every rank applies its slice, the 6*Nc partial is all-reduced inside that kernel over NVLink peer mailboxes") if explicit else
                             ("Schur complement + SCHUR_JACOBI PCG; implicit operator (two passes over the observations per CG iteration: the explicit "
                              "tiles would be a larger stream than the observations); multi-GPU: the 6*Nc partial is all-reduced inside the fused PCG "
                              "kernel over NVLink peer mailboxes"),
            "step": "one LM iteration, fixed count (tolerances 0); solves of at most 6 LM iterations, each restarting from the same initial values",
            "l2": ("inputs larger than L2, no flush: ~1.2 GB of observation planes per linearisation and 479 MB of tiles (330 MB as block-sparse panels) per CG iteration vs 126 MB L2; "
                   "inside a PCG solve 80 MB of the tiles are deliberately kept L2-resident (evict_last) across its CG iterations"),
            "parallelism": f"points sharded over {a.gpus} GPU(s), cameras replicated"}


# --------------------------------------------------------------------------- B200 arm
# algorithmic bytes per observation of the kernels bench.py reports on (DESIGN.md "Kernels")
ALG_BYTES = {
    "k_residual_jacobian": 112,      # in 16 uv + 4 idx + 4 tile slack (+ point gather from L2), out 16 res + 48 A + 24 Pbar.  SURVEY 8(d) counts 184 for the
                                     # fully materialised Jacobian: J_t = -J_X is not stored and J_r = A [R Pbar]x Q^T is rebuilt by its readers
    "k_camera_pass": 60,             # factored form: 24 Pbar + 4 pt idx + 32 gathered y_p  (materialised form: 132)
    "k_point_pass": 56,              # 48 A planes + 4 pt idx + 4 cam idx; C x_r rebuilt as A((X-t) x R Jr x_r) (per-camera operand and the point
                                     # from L1/L2; ~26 B/obs of Hinv / y_p per point not counted)
}
ALG_BYTES.update({
    "k_linearize_q": 72,             # K1p + K2a: in 16 uv + 8 idx, out 48 (six point-order A planes) per observation; + 96 B per point (Hpp, gp) below
    "k_camera_blocks": 88,           # K2b: res 16 + A 48 + Pbar 24 in (coalesced planes); 224 B per tile + 216 B per camera out (not counted)
    "k_schur_diag": 76,              # A 48 + Pbar 24 + pt idx 4 in; + 64 B Hinv gather per observation from L2 (not counted)
})
ALG_BYTES_PER_POINT = {"k_linearize_q": 96}
STAGE = {"k_residual_jacobian": 0, "k_camera_pass": 7, "k_point_pass": 6, "k_linearize_q": 5, "k_camera_blocks": 9, "k_schur_diag": 4}
TILE_BYTES = 384 * 384 * 8           # one dense tile of the explicit reduced camera system


def run_b200(a):
    import torch
    import torch.distributed as dist
    from steviapp_b200 import SBASolver, synthetic, comm_unique_id

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl b200 needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    t0 = time.perf_counter()
    scene = synthetic.make_scene(a.shape, mixed=a.mixed)
    if a.free_intrinsics:   # every block but B free (B's Jacobian is identically zero), start 0.3 % / 2 px / 0.003 off the truth
        intr = scene["intrinsics"].copy()
        intr[0, 0] *= 1.003; intr[0, 1] += 2.0; intr[0, 2] -= 1.5; intr[0, 3] += 0.003; intr[0, 6] += 2e-5
        scene["intrinsics"] = intr
        scene["intr_fixed_mask"] = np.array([0x10], np.uint8)
    full_obs = scene["n_obs"]
    if world > 1:
        scene = synthetic.shard_scene(scene, rank, world)
    t_gen = time.perf_counter() - t0

    s = SBASolver(device=local, pcg_rel_tol=a.pcg_tol, pcg_max_iters=a.pcg_max_iters,
                  pcg_check_every=a.pcg_check_every, linear_solver=a.linear_solver, **FIXED)
    if world > 1:
        uid = [comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        s.comm_init(rank, world, uid[0])
    t0 = time.perf_counter()
    raxis, t, pts = synthetic.load_into(s, scene)
    s.finalize()
    t_ingest = time.perf_counter() - t0
    init = (raxis.copy(), t.copy(), pts.copy())
    intr0 = s.intrinsics.copy() if a.free_intrinsics else None

    def reset():
        raxis[:] = init[0]; t[:] = init[1]; pts[:] = init[2]
        if intr0 is not None:
            s.intrinsics[:] = intr0
        s.upload_parameters()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def maxr(x):
        if world == 1:
            return x
        v = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        return float(v.item())

    # ---- warm-up (untimed): W LM iterations, resident
    reset()
    if a.warmup > 0:
        s.solve(max_iterations=a.warmup, copy_params=False)

    # ---- `value`: K LM iterations, inputs resident in HBM, device time (CUDA events on the solver's stream).  The LM runs in
    # chunks of at most LM_CHUNK iterations, every chunk restarting from the same initial values (a longer uninterrupted run
    # converges to machine precision and stops taking steps); the parameter re-upload between chunks is outside the device time.
    r = None
    first_history = []
    ms_total = 0.0
    wall = 0.0
    done = 0
    clk = ClockSampler(local)   # samples nvidia-smi during BOTH timed regions (value and e2e)
    clk.__enter__()
    while done < a.steps:
        n = min(LM_CHUNK, a.steps - done)
        reset()
        barrier()
        w0 = time.perf_counter()
        rc = s.solve(max_iterations=n, copy_params=False)
        barrier()
        wall += time.perf_counter() - w0
        assert rc["iterations"] == n, rc
        ms_total += maxr(rc["ms_total"])
        done += n
        if r is None:
            r = rc
            first_history = list(rc["cost_history"])
        else:   # accumulate what the report uses
            for k in ("iterations", "pcg_iterations_total", "kernel_launches", "ms_total", "ms_linearize", "ms_schur", "ms_pcg", "ms_update"):
                r[k] += rc[k]
            r["final_cost"] = rc["final_cost"]
    wall = maxr(wall)
    value = a.steps / (ms_total * 1e-3)
    # cost after the FIRST solve (min(LM_CHUNK, K) LM iterations from the initial values): the number the reference arm reports too
    final_cost = first_history[-1]

    # ---- `e2e`: the same K LM iterations through the public call with HOST buffers: every step uploads the
    # parameter blocks (pinned staging), runs one LM iteration and reads the updated blocks back.
    raxis[:] = init[0]; t[:] = init[1]; pts[:] = init[2]
    if intr0 is not None:
        s.intrinsics[:] = intr0
    barrier()
    w0 = time.perf_counter()
    e2e_launches = 0
    for _ in range(a.steps):
        rr = s.solve(max_iterations=1, copy_params=True)
        e2e_launches += rr["kernel_launches"]
    barrier()
    e2e_wall = maxr(time.perf_counter() - w0)
    intr_final = s.intrinsics.copy() if a.free_intrinsics else None
    clk.__exit__()
    nbytes = 8 * (raxis.size + t.size + pts.size)
    e2e = {"value": a.steps / e2e_wall, "unit": UNIT, "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": nbytes,
           "protocol": "K calls of sba_solve(max_iterations=1, copy_params=1); host wall clock between device synchronisations"}

    # ---- kernel-level numbers (rank 0's shard): CUDA events inside the library around back-to-back launches
    reset()
    n_obs_local = s.sizes()["n_obs"]
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_kind = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    kern = {}
    n_pts_local = s.sizes()["n_pts"]
    for name, stage in STAGE.items():
        ms = s.time_stage(stage, 20)
        nb = ALG_BYTES[name] * n_obs_local + ALG_BYTES_PER_POINT.get(name, 0) * n_pts_local
        kern[name] = {"ms": ms, "GBps": nb / (ms * 1e-3) / 1e9, "Mobs_per_s": n_obs_local / ms / 1e3,
                      "alg_bytes_per_launch": nb, "alg_bytes_per_obs": ALG_BYTES[name]}
        if name in ALG_BYTES_PER_POINT:
            kern[name]["alg_bytes_per_point"] = ALG_BYTES_PER_POINT[name]
    pcg_per_step = r["pcg_iterations_total"] / max(r["iterations"], 1)
    ex = s.explicit_info()
    shard = s.explicit_shard()
    if ex["enabled"]:
        # the PCG applies S from the dense tiles: ONE read of every stored scalar per application (symmetric storage);
        # multi-GPU: this rank's slice of the tiles
        ms = s.time_stage(10, 20)
        nbytes = shard["n_own"] * TILE_BYTES
        kern["k_explicit_matvec"] = {"ms": ms, "GBps": nbytes / (ms * 1e-3) / 1e9, "alg_bytes_per_launch": nbytes,
                                     "alg_bytes_per_tile": TILE_BYTES, "tiles": shard["n_own"], "tiles_all_ranks": ex["n_tiles"],
                                     "note": "slice < 126 MB L2: served from L2 after the first application" if nbytes < 120e6 else "streamed from HBM"}
        ms = s.time_stage(11, 10)
        nbytes = ex["n_contributions"] * (12 + 2 * 96) + ex["n_blocks"] * (288 + 16)   # index + write + read of the compact contributions, block out
        kern["explicit_build"] = {"ms": ms, "GBps": nbytes / (ms * 1e-3) / 1e9, "alg_bytes_per_launch": nbytes,
                                  "contributions": ex["n_contributions"], "blocks": ex["n_blocks"]}
        share = {"k_explicit_matvec": pcg_per_step * kern["k_explicit_matvec"]["ms"], "explicit_build": kern["explicit_build"]["ms"],
                 "k_residual_jacobian": kern["k_residual_jacobian"]["ms"]}
        if shard["persistent_pcg"]:
            # the PCG of an LM iteration is ONE launch of the persistent kernel k_pcg_solve (mat-vec, folds, exchange and vector
            # updates of every CG iteration): duration = the PCG phase of the TIMED region above (CUDA events on the solver's
            # stream around that single launch), algorithmic bytes = one read of this rank's tiles per CG iteration
            ms = r["ms_pcg"] / a.steps
            nbytes = pcg_per_step * shard["n_own"] * TILE_BYTES
            kern["k_pcg_solve"] = {"ms": ms, "GBps": nbytes / (ms * 1e-3) / 1e9, "alg_bytes_per_launch": nbytes,
                                   "cg_iterations_per_launch": pcg_per_step, "us_per_cg_iteration": ms * 1e3 / max(pcg_per_step, 1),
                                   "alg_bytes_per_cg_iteration": shard["n_own"] * TILE_BYTES, "tiles": shard["n_own"],
                                   "note": ("slice of the tiles < 126 MB L2: re-read from L2 after the first CG iteration, the HBM roofline does not bound it"
                                            if shard["n_own"] * TILE_BYTES < 120e6 else "tiles streamed from HBM every CG iteration; the leading 80 MB are loaded evict_last and re-read from L2, so the "
                                                 "algorithmic rate can exceed the HBM peak (roofline.traffic = DRAM bytes measured by ncu)")}
            sp = s.explicit_sparse_info()
            if sp["enabled"]:
                # block-sparse panels: the kernel streams only the 6x6 blocks of camera pairs that share a point.  Algorithmic
                # bytes = the values of those blocks (288 B each), once per CG iteration; the stream adds the 16-bit column
                # maps, headers and 128-byte padding (stream_bytes).  dense_equivalent = the rate the dense tiles would need.
                per_it = 288 * sp["n_blocks"]
                kern["k_pcg_solve"].update({
                    "GBps": pcg_per_step * per_it / (ms * 1e-3) / 1e9, "alg_bytes_per_launch": pcg_per_step * per_it,
                    "alg_bytes_per_cg_iteration": per_it, "stream_bytes_per_cg_iteration": sp["stream_bytes"],
                    "blocks": sp["n_blocks"], "panels": sp["n_panels"], "chunks": sp["n_chunks"], "tile_format": "block-sparse panels",
                    "dense_equivalent_GBps": nbytes / (ms * 1e-3) / 1e9,
                    "note": "block-sparse panels streamed from HBM every CG iteration through a byte ring of bulk copies; the leading 80 MB are loaded "
                            "evict_last and re-read from L2; dense_equivalent_GBps = dense tile bytes / time, for comparison with the dense-tile runs"})
            kern["k_explicit_matvec"]["note"] = "stand-alone tile mat-vec kernel (operator hook / fallback), not launched by the PCG"
            share = {"k_pcg_solve": ms, "explicit_build": kern["explicit_build"]["ms"], "k_residual_jacobian": kern["k_residual_jacobian"]["ms"]}
    else:
        # implicit operator: the two passes run once per CG iteration, everything else once per linearisation / LM iteration
        share = {k: (pcg_per_step if k in ("k_camera_pass", "k_point_pass") else 1.0) * v["ms"] for k, v in kern.items()}
    dom = max(share, key=share.get)
    traffic = None
    traffic_source = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic_source = "profiles/traffic.json: " + str(tj.get("_source", "ncu --set full capture (dram__bytes_read.sum + dram__bytes_write.sum)")) + "; not measured in this run"
        traffic = tj.get(
            dom + "_sparse" if kern[dom].get("tile_format") == "block-sparse panels" else dom)
        if isinstance(traffic, dict):   # persistent kernel: ncu measured one launch of 114 CG iterations; per launch of THIS run
            traffic = traffic["bytes_per_cg_iteration"] * pcg_per_step * (shard["n_own"] / max(ex["n_tiles"], 1))
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": dom, "achieved": kern[dom]["GBps"], "peak": peak, "unit": "GB/s",
                "frac": kern[dom]["GBps"] / peak, "traffic": traffic, "traffic_source": traffic_source, "peak_source": peak_kind,
                "alg_bytes_per_launch": kern[dom]["alg_bytes_per_launch"], "obs_per_launch": n_obs_local,
                "share_of_step": share[dom] / (ms_total / a.steps),
                "kernels": {k: dict(v, frac=v["GBps"] / peak) for k, v in kern.items()}}

    # ---- labelled second figure: the same K LM iterations as an INEXACT Newton method (CG stopped at relative residual eta, as
    # Ceres runs ITERATIVE_SCHUR with its default eta = 0.1).  Never the headline: the headline solves to pcg_rel_tol.
    inexact = None
    if a.eta > 0.0 and a.linear_solver == 0:
        s.set_options(pcg_eta=a.eta)
        ms_i, done_i, pcg_i, hist_i = 0.0, 0, 0, None
        while done_i < a.steps:
            nn = min(LM_CHUNK, a.steps - done_i)
            reset()
            barrier()
            rc = s.solve(max_iterations=nn, copy_params=False)
            barrier()
            ms_i += maxr(rc["ms_total"]); done_i += rc["iterations"]; pcg_i += rc["pcg_iterations_total"]
            if hist_i is None:
                hist_i = list(rc["cost_history"])
            if rc["iterations"] < nn:
                break
        s.set_options(pcg_eta=0.0)
        inexact = {"pcg_eta": a.eta, "value": done_i / (ms_i * 1e-3), "unit": UNIT, "ms_per_step": ms_i / max(done_i, 1),
                   "pcg_iterations_per_step": pcg_i / max(done_i, 1), "cost_history_first_solve": hist_i,
                   "headline_cost_history_first_solve": first_history,
                   "note": "inexact Newton (forcing term eta): a throughput setting of the same solver, reported beside the headline, never instead of it"}

    # ---- BASELINE config 4: batched multi-view DLT of --dlt-points million points (the scene's views replicated)
    dlt = None
    if a.dlt_points > 0 and world == 1:
        truth = synthetic.make_scene(a.shape, mixed=a.mixed)
        raxis[:] = truth["raxis_true"]; t[:] = truth["t_true"]
        s.upload_parameters()
        order = np.argsort(truth["obs_pt"], kind="stable")
        cam1 = truth["obs_cam"][order]; uv1 = truth["obs_uv"][order]
        cnt1 = np.bincount(truth["obs_pt"], minlength=truth["n_pts"])
        repn = int(np.ceil(a.dlt_points * 1e6 / truth["n_pts"]))
        view_cam = np.tile(cam1, repn); view_uv = np.tile(uv1, (repn, 1))
        view_off = np.concatenate([[0], np.cumsum(np.tile(cnt1, repn))]).astype(np.int32)
        npt, nv = view_off.size - 1, int(view_off[-1])
        best = None
        for _ in range(3):
            tw = time.perf_counter()
            x, okf, ms = s.triangulate(view_off, view_cam, view_uv)
            wall_d = time.perf_counter() - tw
            best = ms if best is None else min(best, ms)
        survey_bytes = 72 * nv + 24 * npt          # 72 B per view (pose + ray) + 24 B per point out = 384 B per point at 5 views
        stream_bytes = nv * (4 + 16) + npt * (4 + 24 + 1)
        dlt = {"n_points": npt, "n_views": nv, "kernel_ms": best, "Gpoints_per_s": npt / best / 1e6,
               "alg_GBps_survey_bytes": survey_bytes / best / 1e6, "frac_of_hbm_peak_survey_bytes": survey_bytes / best / 1e6 / peak,
               "streamed_GBps": stream_bytes / best / 1e6, "e2e_wall_s_with_host_copies": wall_d, "ok_fraction": float(okf.mean()),
               "max_abs_err_vs_truth_first_copy": float(np.abs(x[:truth["n_pts"]] - truth["pts_true"]).max()),
               "note": "camera records (64 B: quaternion + Rt t) are gathered from L1/L2, only pixels + camera indices stream from HBM: the SURVEY byte count (poses and rays per view) is an upper bound of the traffic"}
        reset()

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline and not a.free_intrinsics and not a.mixed:
        # bounded sample: 3 LM iterations of the full scene (the parity tests' count), no warm-up; the cost history is held
        # against the GPU's below
        cpu = cpu_leg(a.shape, 3 if scene["n_obs"] < 10_000_000 else 1, 0, a.pcg_tol, a.pcg_max_iters, scene=scene, budget_s=60.0)
        gh = first_history
        ch = cpu["cost_history_first_solve"]
        m = min(len(gh), len(ch))
        cpu["gpu_cost_history_first_solve"] = gh[:m]
        cpu["max_rel_cost_diff_vs_gpu"] = max(abs(x - y) / y for x, y in zip(gh[:m], ch[:m]))

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": bench_config(a, bool(ex["enabled"])),
            "clocks": clk.summary(), "e2e": e2e, "gpu_launches": r["kernel_launches"],
            "roofline": roofline, "cpu_baseline": cpu,
            "residual_jacobian_mobs_per_s": kern["k_residual_jacobian"]["Mobs_per_s"] * (full_obs / n_obs_local if world > 1 else 1.0),
            "pcg_iterations_per_step": pcg_per_step, "pcg_allreduce": ("peer mailboxes (fused kernel)" if shard["peer_exchange"] else
                                                                        ("nccl" if world > 1 else "none")), "final_cost": final_cost, "final_cost_after_lm_iterations": len(first_history) - 1, "cost_history_first_solve": first_history, "initial_cost": r["initial_cost"],
            "phase_ms": {k: r[k] for k in ("ms_linearize", "ms_schur", "ms_pcg", "ms_update")},
            "wall_ms_per_step": wall * 1e3 / a.steps, "ingest_s": t_ingest, "scene_gen_s": t_gen,
            "inexact_newton": inexact, "dlt": dlt,
        }
        if a.free_intrinsics:
            line["intrinsics"] = {"initial": intr0[0].tolist(), "after_the_e2e_steps": intr_final[0].tolist(), "truth": synthetic.DEFAULT_INTRINSICS.tolist(),
                                  "shared": s.shared_info()["n_ext"]}
        print(json.dumps(line), flush=True)
    s.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
