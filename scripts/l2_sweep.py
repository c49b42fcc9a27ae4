"""Sweep of the L2-resident share of the tiles in the persistent PCG kernel (SBA_PS_L2_MB is read at the first launch of a process,
so every value runs in its own process).  Usage: l2_sweep.py [shape]"""
import json
import os
import subprocess
import sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
code = f"""
import sys; sys.path.insert(0, {root!r})
from steviapp_b200 import SBASolver, synthetic
sc = synthetic.make_scene({shape!r})
s = SBASolver(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0)
synthetic.load_into(s, sc)
r = s.solve(max_iterations=4)
n = sum(r['pcg_iters_history'])
print('%.2f us/CG iteration, %.3f ms/LM iteration, pcg %.2f ms, final %.10g' % (r['ms_pcg'] / n * 1e3, r['ms_total'] / 4, r['ms_pcg'], r['final_cost']))
"""
for mb in (sys.argv[2:] or ["0", "40", "60", "80", "100", "120"]):
    env = dict(os.environ, SBA_PS_L2_MB=mb)
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    print(f"SBA_PS_L2_MB={mb:>4s}: {out.stdout.strip() or out.stderr[-300:]}", flush=True)
