"""Device time of the explicit build (contributions + assembly) on a synthetic scene.  Usage: build_time.py [shape] [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from steviapp_b200 import SBASolver, synthetic

shape = sys.argv[1] if len(sys.argv) > 1 else "venice"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
s = SBASolver()
synthetic.load_into(s, synthetic.make_scene(shape)); s.finalize()
print(f"explicit build {s.time_stage(11, reps) * 1e3:.1f} us")
