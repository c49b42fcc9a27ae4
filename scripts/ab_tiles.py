"""A/B of the prefetching variants of the camera-order tile kernels on the Venice shape: stage 9 = K2b camera blocks (+ reduce,
finish), stage 4 = Schur diagonal (+ reduce, finish, inversion), stage 0 / 1 = K1 / K1c."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

sc = synthetic.make_scene(sys.argv[1] if len(sys.argv) > 1 else "venice")
s = SBASolver()
synthetic.load_into(s, sc)
s.finalize()
s.solve(max_iterations=1, function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)   # fills every plane the stages read
for rep in range(2):
    for pipe in ("0", "2", "3"):
        for k in ("SBA_K2B_PIPE", "SBA_SD_PIPE", "SBA_K1_PIPE"):
            os.environ[k] = pipe
        print("pipe", pipe, "K2b ms %.4f" % s.time_stage(9, 50), "schur_diag(+invert) ms %.4f" % s.time_stage(4, 50),
              "K1 ms %.4f" % s.time_stage(0, 50), "K1c ms %.4f" % s.time_stage(1, 50), flush=True)
for pipe in ("0", "2"):
    for k in ("SBA_K2B_PIPE", "SBA_SD_PIPE", "SBA_K1_PIPE"):
        os.environ[k] = pipe
    s2 = SBASolver()
    synthetic.load_into(s2, sc)
    r = s2.solve(max_iterations=3, function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
    print("pipe", pipe, [repr(c) for c in r["cost_history"]], r["pcg_iters_history"][:4])
    s2.close()
