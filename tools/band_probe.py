"""dev probe: volume build of a row band of the 4096x4096x72 workload (what one rank of an N-GPU job runs),
   cold and warm L2, CUDA events.   python tools/band_probe.py [builds]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from path_planner_b200 import capi
from path_planner_b200.synth import make_costmap
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10
cost = make_costmap(4096, 4096, seed=1, pockets=[(3000, 1000, 6)])
h = capi.Handle(0); h.costmap_upload(cost); h.footprint_from_boundary(0.05, [0.64, 0.46], True)
for rows in [int(r) for r in os.environ.get('BANDS', '16,32,64,128,256,512,1024,2048,4096').split(',')]:
    y0 = 1024 if rows <= 2048 else 0
    h.se2_volume_set_rows(y0, y0 + rows)
    for _ in range(4):
        h.se2_volume_build(72)
    h.sync()
    cold, warm = [], []
    for _ in range(n):
        h.l2_flush(); h.timer_start(); h.se2_volume_build(72); cold.append(h.timer_stop())
    for _ in range(n):
        h.timer_start(); h.se2_volume_build(72); warm.append(h.timer_stop())
    h.timer_start()
    for _ in range(n):
        h.se2_volume_build(72)
    b2b = h.timer_stop() / n
    print(f"rows {rows:5d}: cold {np.mean(cold)*1e3:8.1f} us (min {np.min(cold)*1e3:8.1f})  warm {np.mean(warm)*1e3:8.1f}  back-to-back {b2b*1e3:8.1f}  per-row-of-16 {np.mean(cold)*1e3/(rows/16):6.2f}", flush=True)
