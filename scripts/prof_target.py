"""Short profiling target: two LM iterations on a synthetic scene.  Usage: prof_target.py [shape] [pcg_max_iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
cap = int(sys.argv[2]) if len(sys.argv) > 2 else 500
sc = synthetic.make_scene(shape)
s = SBASolver(pcg_check_every=8, pcg_max_iters=cap, function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0,
              linear_solver=int(os.environ.get("SBA_PROF_LINEAR_SOLVER", "0")))
synthetic.load_into(s, sc)
r = s.solve(max_iterations=int(os.environ.get("SBA_PROF_ITERS", "2")))
print(r["iterations"], r["final_cost"], r["kernel_launches"], r["pcg_iters_history"])
