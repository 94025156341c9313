"""dev probe: build time (CUDA events, third and later builds, i.e. the thorough plan) per footprint and map size."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from path_planner_b200 import capi
from path_planner_b200.synth import make_costmap
CASES = [(4096, 72, [0.64, 0.46], True, "rect hole (P=48, symmetric)"), (4096, 72, [0.64, 0.46], False, "rect solid (P=165, symmetric)"),
         (4096, 72, [-0.2, -0.25, 0.6, 0.25], True, "box hole (P=56, asymmetric)"), (4096, 72, [0.35], True, "circle hole"),
         (8192, 144, [0.64, 0.46], True, "rect hole, 8192^2 x 144")]
for (W, nt, b, hole, name) in CASES:
    cost = make_costmap(W, W, seed=1)
    h = capi.Handle(0); h.costmap_upload(cost); h.footprint_from_boundary(0.05, b, hole)
    for _ in range(4):
        h.se2_volume_build(nt)
    h.sync()
    ms = 0.0
    for _ in range(10):
        h.l2_flush(); h.timer_start(); h.se2_volume_build(nt); ms += h.timer_stop()
    st = h.se2_volume_stats()
    print(f"{name:34s} {ms / 10:7.3f} ms  {W * W * nt / (ms / 10 * 1e-3):.3e} cells/s  loads/cell {st['samples_per_cell']:.2f}  generic tile-bins {st['generic_tiles']}")
    h.close()
