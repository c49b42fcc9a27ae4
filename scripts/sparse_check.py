"""Dense tiles vs block-sparse panels in the persistent PCG kernel: same LM trajectory, time per CG iteration.
Usage: sparse_check.py [shape ...]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

fixed = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
for shape in (sys.argv[1:] or ["ladybug", "venice"]):
    sc = synthetic.make_scene(shape)
    out = {}
    for sp in ("1", "0"):
        os.environ["SBA_PS_SPARSE"] = sp
        s = SBASolver(pcg_check_every=8, **fixed)
        synthetic.load_into(s, sc)
        s.finalize()
        assert s.explicit_shard()["persistent_pcg"]
        r = s.solve(max_iterations=4)
        out[sp] = r
        s.close()
        its = sum(r["pcg_iters_history"])
        print(json.dumps({"shape": shape, "sparse": sp, "cost": r["cost_history"], "pcg": r["pcg_iters_history"],
                          "ms_total": r["ms_total"], "ms_pcg": r["ms_pcg"], "us_per_cg_it": 1e3 * r["ms_pcg"] / max(its, 1)}))
    rel = max(abs(a - b) / max(b, 1.0) for a, b in zip(out["1"]["cost_history"], out["0"]["cost_history"]))
    print(json.dumps({"shape": shape, "max_rel_cost_diff": rel}))
    assert rel < 1e-8
