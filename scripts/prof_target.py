"""Short profiling target: one linearisation + a few operator applications on a synthetic scene."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
sc = synthetic.make_scene(shape)
s = SBASolver(pcg_check_every=8, pcg_max_iters=6, function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
synthetic.load_into(s, sc)
r = s.solve(max_iterations=2)
print(r["iterations"], r["final_cost"], r["kernel_launches"])
