import sys, os
sys.path.insert(0, "/root/repo")
from steviapp_b200 import SBASolver, synthetic
sc = synthetic.make_scene("venice")
s = SBASolver()
synthetic.load_into(s, sc)
s.finalize()
print("variant", os.environ.get("SBA_PP_VARIANT", "default"), "k_linearize_q ms", s.time_stage(5, 20), "K1 ms", s.time_stage(0, 20), "pp", s.time_stage(6, 20))
r = s.solve(max_iterations=2, function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
print(r["cost_history"])
