import sys, numpy as np
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from path_planner_b200 import capi
from path_planner_b200.synth import make_costmap
cost=make_costmap(4096,4096,seed=1)
h=capi.Handle(0); h.costmap_upload(cost); h.footprint_from_boundary(0.05,[0.64,0.46],True)
for nb in (72,36,18,8,4,1):
    for _ in range(3): h.se2_volume_build(72,0.5,0.5,0.0,0,nb)
    h.sync(); ts=[]
    for _ in range(5):
        h.l2_flush(); h.timer_start(); h.se2_volume_build(72,0.5,0.5,0.0,0,nb); ts.append(h.timer_stop())
    print(nb, 'bins: %.3f ms'%np.mean(ts), h.se2_volume_stats())
