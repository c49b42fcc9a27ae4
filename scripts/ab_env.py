"""Generic A/B over one environment knob on the Venice shape: per-LM-iteration phase times of a 6-iteration solve (after a warm-up
solve) and the cost history.  Usage: ab_env.py VAR v1,v2,... [shape]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

var, vals = sys.argv[1], sys.argv[2].split(",")
sc = synthetic.make_scene(sys.argv[3] if len(sys.argv) > 3 else "venice")
fixed = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
for rep in range(2):
    for v in vals:
        os.environ[var] = v
        s = SBASolver(pcg_check_every=8, **fixed)
        synthetic.load_into(s, sc)
        s.finalize()
        s.solve(max_iterations=1, copy_params=False)
        r = s.solve(max_iterations=6, copy_params=False)
        n = r["iterations"]
        print(var, v, "ms/it total %.3f lin %.3f schur %.3f pcg %.3f update %.3f" % (r["ms_total"] / n, r["ms_linearize"] / n, r["ms_schur"] / n,
              r["ms_pcg"] / n, r["ms_update"] / n), "final", repr(r["final_cost"]), flush=True)
        s.close()
