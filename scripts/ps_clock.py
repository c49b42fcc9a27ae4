"""A/B of the PCG paths on one GPU: persistent kernel (with its phase clock, SBA_PS_CLOCK=1) vs mat-vec + fused kernel.
Usage: ps_clock.py [shape] [lm_iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
sc = synthetic.make_scene(shape)
for persistent, clock in (("1", ""), ("0", ""), ("1", "1"), ("1", "")):
    os.environ["SBA_PCG_PERSISTENT"] = persistent
    if clock:
        os.environ["SBA_PS_CLOCK"] = clock
    else:
        os.environ.pop("SBA_PS_CLOCK", None)
    s = SBASolver(function_tolerance=0.0, parameter_tolerance=0.0, gradient_tolerance=0.0, pcg_check_every=8)
    synthetic.load_into(s, sc)
    r = s.solve(max_iterations=iters)
    n = sum(r["pcg_iters_history"])
    print(f"persistent={persistent} clock={clock or 0}: pcg iters {r['pcg_iters_history']}  ms_pcg {r['ms_pcg']:.2f}  us/iter {r['ms_pcg'] / n * 1e3:.1f}  ms_total {r['ms_total']:.2f}", flush=True)
    s.close()
