"""Short profiling target for K5: batched DLT of ~2 M points (the Venice-shape views, mixed plane / spherical cameras, replicated twice)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from steviapp_b200 import SBASolver, synthetic

sc = synthetic.make_scene("venice", mixed=True)
sc["raxis"], sc["t"] = sc["raxis_true"].copy(), sc["t_true"].copy()
order = np.argsort(sc["obs_pt"], kind="stable")
cnt = np.bincount(sc["obs_pt"], minlength=sc["n_pts"])
view_cam = np.tile(sc["obs_cam"][order], 2); view_uv = np.tile(sc["obs_uv"][order], (2, 1))
view_off = np.concatenate([[0], np.cumsum(np.tile(cnt, 2))]).astype(np.int32)
s = SBASolver()
synthetic.load_into(s, sc)
x, ok, ms = s.triangulate(view_off, view_cam, view_uv)
print(view_off.size - 1, "points", ms, "ms", float(ok.mean()))
for ex in os.environ.get("SBA_DLT_EXP_LIST", "").split(","):   # timing experiments (results are garbage for ex != 0)
    if ex:
        os.environ["SBA_DLT_EXP"] = ex
        t = min(s.triangulate(view_off, view_cam, view_uv)[2] for _ in range(3))
        print("experiment", ex, "kernel ms %.4f" % t, "per point ns %.3f" % (1e6 * t / (view_off.size - 1)))
