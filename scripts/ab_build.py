"""A/B of the explicit-build variants (SBA_EX_ASM2) on the Venice shape: stage 11 = contributions + assembly."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

sc = synthetic.make_scene(sys.argv[1] if len(sys.argv) > 1 else "venice")
s = SBASolver()
synthetic.load_into(s, sc)
s.finalize()
s.solve(max_iterations=1, function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
for rep in range(2):
    for v in ("0", "1", "2"):
        os.environ["SBA_EX_ASM2"] = v
        print("asm2", v, "build ms %.4f" % s.time_stage(11, 30), flush=True)
for v in ("0", "1", "2"):
    os.environ["SBA_EX_ASM2"] = v
    s2 = SBASolver()
    synthetic.load_into(s2, sc)
    r = s2.solve(max_iterations=3, function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
    print("asm2", v, [repr(c) for c in r["cost_history"]], r["pcg_iters_history"][:4], "ms_schur/it %.3f" % (r["ms_schur"] / 3))
    s2.close()
