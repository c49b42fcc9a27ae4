"""One rank's share of the city shape (BASELINE config 5: 10 000 cameras / 10 M points / 50 M observations over 8 GPUs) on ONE GPU, implicit
operator, a capped PCG: the kernel split of the path the 8-GPU run takes (minus its peer exchange).  Usage: prof_city_shard.py [pcg_cap]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

cap = int(sys.argv[1]) if len(sys.argv) > 1 else 24
t0 = time.time()
sc = synthetic.shard_scene(synthetic.make_scene("city"), 0, 8)
print("scene + shard %.1f s" % (time.time() - t0), sc["n_cams"], sc["n_pts"], sc["n_obs"], flush=True)
s = SBASolver(pcg_check_every=8, pcg_max_iters=cap, operator_mode=2, function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
synthetic.load_into(s, sc)
r = s.solve(max_iterations=2)
print(r["iterations"], r["pcg_iters_history"][:3], "ms: pcg %.3f schur %.3f lin %.3f update %.3f" % (r["ms_pcg"], r["ms_schur"], r["ms_linearize"], r["ms_update"]),
      "us per CG iteration %.1f" % (1e3 * r["ms_pcg"] / max(sum(r["pcg_iters_history"]), 1)))
for stage, name in ((6, "point pass"), (7, "camera pass"), (12, "fused vector kernel")):
    try:
        print(name, "ms %.4f" % s.time_stage(stage, 20))
    except Exception as e:
        print(name, "n/a", str(e)[:80])
