"""A/B of the K1 variants (SBA_K1_PIPE) on the Venice shape: stage 0 = K1 (residual + Jacobian), stage 1 = K1c (cost only)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

sc = synthetic.make_scene(sys.argv[1] if len(sys.argv) > 1 else "venice")
s = SBASolver()
synthetic.load_into(s, sc)
s.finalize()
for rep in range(2):
    for pipe, minb in (("0", "4"), ("0", "3"), ("1", "4"), ("1", "3"), ("1", "2")):
        os.environ["SBA_K1_PIPE"] = pipe; os.environ["SBA_K1_MINB"] = minb
        print("pipe", pipe, "minb", minb, "K1 ms %.4f" % s.time_stage(0, 50), "K1c ms %.4f" % s.time_stage(1, 50), flush=True)
os.environ["SBA_K1_PIPE"] = "1"; os.environ.pop("SBA_K1_MINB")
r = s.solve(max_iterations=2, function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
print(r["cost_history"])
