"""Device time of the PCG iteration pieces on a synthetic scene (explicit system, one GPU). Usage: pcg_times.py [shape] [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
sc = synthetic.make_scene(shape)
s = SBASolver(pcg_check_every=8)
synthetic.load_into(s, sc); s.finalize()
for st, name in ((10, "tile mat-vec"), (12, "fused PCG kernel alone"), (13, "mat-vec + fused kernel")):
    print(f"{name:28s} {s.time_stage(st, reps) * 1e3:8.2f} us")
