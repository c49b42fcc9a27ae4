"""Sensitivity of the final cost after a fixed number of LM iterations to the PCG tolerance (Venice shape).
The parity bar is 1e-8 relative on that cost; this shows which tolerances meet it.  Usage: tol_study.py [shape] [iters]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 9
sc = synthetic.make_scene(shape)
fixed = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
rows = []
for tol in (1e-13, 1e-10, 1e-8, 1e-6, 1e-4, 1e-2):
    with SBASolver(pcg_check_every=8, pcg_rel_tol=tol, pcg_max_iters=1000) as s:
        synthetic.load_into(s, sc)
        r = s.solve(max_iterations=iters, **fixed)
    rows.append({"tol": tol, "final_cost": r["final_cost"], "pcg_total": r["pcg_iterations_total"], "ms_total": r["ms_total"],
                 "cost_history": list(r["cost_history"][:iters + 1]), "successful": r["successful_steps"]})
ref = rows[0]["final_cost"]
for r in rows:
    r["rel_to_tightest"] = abs(r["final_cost"] - ref) / ref
    print(f"tol {r['tol']:.0e}  final {r['final_cost']:.12e}  rel diff {r['rel_to_tightest']:.2e}  pcg its {r['pcg_total']:5d}  ms {r['ms_total']:.1f}  ok steps {r['successful']}")
json.dump(rows, open(os.path.join("gpurun_out", f"tol_study_{shape}.json"), "w"), indent=1)
