"""dev probe: the headline volume build alone (CUDA events, cold L2), short enough to run under ncu.
   python tools/k1_probe.py [rect|box|solid] [builds]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from path_planner_b200 import capi
from path_planner_b200.synth import make_costmap
kind = sys.argv[1] if len(sys.argv) > 1 else "rect"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 10
fp = {"rect": ([0.64, 0.46], True), "box": ([-0.2, -0.25, 0.6, 0.25], True), "solid": ([0.64, 0.46], False)}[kind]
cost = make_costmap(4096, 4096, seed=1, pockets=[(3000, 1000, 6)])
h = capi.Handle(0); h.costmap_upload(cost); h.footprint_from_boundary(0.05, fp[0], fp[1])
if os.environ.get('ROWS'):
    y0, y1 = map(int, os.environ['ROWS'].split(','))
    h.se2_volume_set_rows(y0, y1)
for _ in range(4):
    h.se2_volume_build(72)
h.sync()
ts = []
for _ in range(n):
    h.l2_flush(); h.timer_start(); h.se2_volume_build(72); ts.append(h.timer_stop())
st = h.se2_volume_stats()
print(f"{kind}: {np.mean(ts):.4f} ms (min {np.min(ts):.4f})  loads/cell {st['samples_per_cell']:.2f}  generic tile-bins {st['generic_tiles']}  tables {os.environ.get('PP_SE2_MAX_TABLES', 'auto')}")
