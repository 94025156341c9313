"""dev probe: heuristic build time at 4096^2 on the bench map (PP_HEUR_DELTA is read once per process)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from path_planner_b200 import capi
cost = bench.make_inputs()
h = capi.Handle(0); h.costmap_upload(cost)
for _ in range(2):
    h.heuristic_build(bench.GOAL[0], bench.GOAL[1], 2.0, download=False)
h.sync(); h.timer_start()
for _ in range(5):
    h.heuristic_build(bench.GOAL[0], bench.GOAL[1], 2.0, download=False)
ms = h.timer_stop() / 5
print("delta", os.environ.get("PP_HEUR_DELTA", "default"), "ms %.3f" % ms, h.heuristic_stats())
