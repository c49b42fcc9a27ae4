"""Short profiling target for the shared-block path: two LM iterations on the Venice shape with one free intrinsics model."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from steviapp_b200 import SBASolver, synthetic

sc = synthetic.make_scene(sys.argv[1] if len(sys.argv) > 1 else "venice")
sc["intr_fixed_mask"] = np.array([0x10], np.uint8)          # f, pp, k, t free; B constant
s = SBASolver(pcg_check_every=8, function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
synthetic.load_into(s, sc)
r = s.solve(max_iterations=int(os.environ.get("SBA_PROF_ITERS", "2")))
print(r["iterations"], r["final_cost"], r["kernel_launches"], r["pcg_iters_history"][:4],
      "ms/it schur %.3f pcg %.3f lin %.3f" % (r["ms_schur"] / r["iterations"], r["ms_pcg"] / r["iterations"], r["ms_linearize"] / r["iterations"]))
