"""dev probe: the heuristic field build alone at 4096x4096 (CUDA events), short enough to run under ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from path_planner_b200 import capi
from path_planner_b200.synth import make_costmap
n = int(sys.argv[1]) if len(sys.argv) > 1 else 5
cost = make_costmap(4096, 4096, seed=1, pockets=[(3000, 1000, 6)])
h = capi.Handle(0); h.costmap_upload(cost)
for _ in range(2):
    h.heuristic_build(3000, 1000, 2.0, download=False)
h.sync()
ts = []
for _ in range(n):
    h.timer_start(); h.heuristic_build(3000, 1000, 2.0, download=False); ts.append(h.timer_stop())
print(f"heuristic 4096^2: {np.mean(ts):.3f} ms (min {np.min(ts):.3f})", h.heuristic_stats())
