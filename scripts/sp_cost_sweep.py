"""Run-balance constants of the block-sparse panel schedule (SBA_SP_PANEL_COST / SBA_SP_SEG_COST): time per CG iteration
of the persistent PCG kernel on one scene, two LM iterations per setting.  Usage: sp_cost_sweep.py [shape]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
sc = synthetic.make_scene(shape)
fixed = dict(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
os.environ["SBA_PS_SPARSE"] = "1"
grid = [tuple(int(v) for v in x.split(":")) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else \
    [(panel, seg) for panel in (1024, 2048, 3072, 4608) for seg in (131072, 262144, 393216)]
for g in grid:
    panel, seg = g[0], g[1]
    os.environ["SBA_SP_WMAX_COST"] = str(g[2] if len(g) > 2 else 0)
    if True:
        os.environ["SBA_SP_PANEL_COST"] = str(panel)
        os.environ["SBA_SP_SEG_COST"] = str(seg)
        s = SBASolver(pcg_check_every=8, **fixed)
        synthetic.load_into(s, sc)
        s.finalize()
        s.solve(max_iterations=1, copy_params=False)
        r = s.solve(max_iterations=2, copy_params=False)
        its = sum(r["pcg_iters_history"])
        print(json.dumps({"panel_cost": panel, "seg_cost": seg, "wmax_cost": os.environ["SBA_SP_WMAX_COST"], "us_per_cg_it": 1e3 * r["ms_pcg"] / max(its, 1), "cg_its": its,
                          "ms_total": r["ms_total"]}), flush=True)
        s.close()
