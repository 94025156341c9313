"""dev probe: heuristic field build statistics."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from path_planner_b200 import capi
from path_planner_b200.synth import make_costmap
for (W, goal) in [(4096, (3000, 1000)), (4096, (2048, 2048)), (2048, (100, 100))]:
    cost = make_costmap(W, W, seed=1, pockets=[(goal[0], goal[1], 6)])
    h = capi.Handle(0); h.costmap_upload(cost)
    for _ in range(2): h.heuristic_build(goal[0], goal[1], 2.0, download=False)
    h.sync(); h.timer_start()
    for _ in range(5): h.heuristic_build(goal[0], goal[1], 2.0, download=False)
    ms = h.timer_stop() / 5
    st = h.heuristic_stats()
    nt = ((W + 31) // 32) ** 2
    print(W, goal, "%.2f ms" % ms, st, "tiles", nt, "visits/tile %.2f" % (st["tile_visits"] / nt))
    h.close()
