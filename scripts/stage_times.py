"""Per-stage device times on a synthetic scene (CUDA events inside the library). Usage: stage_times.py [shape] [reps]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
t0 = time.time(); sc = synthetic.make_scene(shape); t_gen = time.time() - t0
s = SBASolver(pcg_check_every=8)
t0 = time.time(); synthetic.load_into(s, sc); s.finalize(); t_fin = time.time() - t0
sz = s.sizes()
names = {0: "K1 residual+jacobian (cam order)", 1: "K1c cost only", 5: "K1p+K2a linearize + point blocks",
         9: "K2b camera blocks", 2: "K1p+K2a+K2b", 4: "schur diag + invert",
         6: "matvec point pass", 7: "matvec camera pass", 3: "matvec S*x"}
No = sz["n_obs"]
bytes_per_obs = {0: 112, 1: 20 + 32, 5: 24 + 48 + 19, 9: 88, 2: 24 + 48 + 19 + 88, 6: 56, 7: 60, 3: 116}
out = {"shape": shape, "sizes": sz, "gen_s": t_gen, "ingest_s": t_fin, "stages": {}}
for st in (0, 1, 5, 9, 2, 4, 6, 7, 3):
    ms = s.time_stage(st, reps)
    gbs = bytes_per_obs.get(st, 0) * No / (ms * 1e-3) / 1e9 if st in bytes_per_obs else None
    out["stages"][names[st]] = {"ms": ms, "GB/s(alg)": gbs, "Mobs/s": No / ms / 1e3}
    print(f"{names[st]:36s} {ms:9.4f} ms  {No / ms / 1e3:10.1f} Mobs/s  {gbs if gbs is None else round(gbs, 1)} GB/s")
fixed = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
r = s.solve(max_iterations=5, **fixed)
keys = ("iterations", "pcg_iterations_total", "initial_cost", "final_cost", "ms_total", "ms_linearize", "ms_schur", "ms_pcg", "ms_update", "ms_h2d", "ms_d2h", "kernel_launches")
out["solve"] = {k: r[k] for k in keys}
out["solve"]["pcg_iters_history"] = r["pcg_iters_history"]
print(json.dumps(out["solve"]))
json.dump(out, open(os.path.join("gpurun_out", f"stage_times_{shape}.json"), "w"), indent=1)
