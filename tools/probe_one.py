"""dev probe: build a given number of heading bins of the 4096x4096 volume a few times (for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from path_planner_b200 import capi
from path_planner_b200.synth import make_costmap
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 1
cost = make_costmap(4096, 4096, seed=1)
h = capi.Handle(0); h.costmap_upload(cost); h.footprint_from_boundary(0.05, [0.64, 0.46], True)
for _ in range(4):
    h.l2_flush(); h.se2_volume_build(72, 0.5, 0.5, 0.0, 0, nb)
h.sync()
