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
