"""dev probe: host planning time (first build) for the benchmark configuration."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from path_planner_b200 import capi
from path_planner_b200.synth import make_costmap
for (W, nt, b, hole) in [(4096, 72, [0.64, 0.46], True), (4096, 72, [0.64, 0.46], False), (2048, 72, [-0.2, -0.25, 0.6, 0.25], True), (8192, 144, [0.64, 0.46], True)]:
    cost = make_costmap(W, W, seed=1)
    h = capi.Handle(0); h.costmap_upload(cost); h.footprint_from_boundary(0.05, b, hole); h.sync()
    t = time.perf_counter(); h.se2_volume_build(nt); h.sync(); t1 = time.perf_counter() - t
    t = time.perf_counter(); h.se2_volume_build(nt); h.sync(); t2 = time.perf_counter() - t
    print(W, nt, b, hole, "first build %.3f s (plan + kernels), second %.4f s" % (t1, t2), h.se2_volume_stats())
    h.close()
