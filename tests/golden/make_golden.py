"""Writes tests/golden/*.json from the CPU oracle (run once here; the fixtures travel to the GPU box).
Usage: python tests/golden/make_golden.py"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py  # noqa: E402
from steviapp_b200 import synthetic  # noqa: E402


def one(name, scene_args, iters):
    sc = synthetic.make_scene(**scene_args)
    p = oracle_py.Problem(sc)
    e = p.evaluate()
    idx = list(range(0, sc["n_obs"], max(1, sc["n_obs"] // 16)))[:16]
    s = oracle_py.Problem(sc).solve(max_iterations=iters, function_tolerance=0.0, parameter_tolerance=0.0,
                                    gradient_tolerance=0.0, num_threads=1)
    out = {
        "scene_args": scene_args, "initial_cost": e["cost"], "sample_obs": idx,
        "residuals": e["residuals"][idx].tolist(), "jac_cam": e["jac_cam"][idx].tolist(),
        "jac_pt": e["jac_pt"][idx].tolist(), "lm_iterations": iters, "final_cost": s["final_cost"],
        "cost_history": s["cost_history"], "pcg_iters_history": s["pcg_iters_history"],
    }
    with open(os.path.join(HERE, name + ".json"), "w") as f:
        json.dump(out, f, indent=1)
    print(name, "initial", e["cost"], "final", s["final_cost"], "iters", s["iterations"])


if __name__ == "__main__":
    one("small_scene", {"shape": "small"}, 5)
    one("small_scene_spherical", {"shape": "small", "spherical": True}, 5)
    one("small_scene_mixed", {"shape": "small", "mixed": True}, 5)
    one("ladybug_scene", {"shape": "ladybug"}, 5)
