"""dev probe: volume build time vs number of brick tables and bins per launch."""
import os, sys, subprocess, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1:
    import numpy as np
    from path_planner_b200 import capi
    from path_planner_b200.synth import make_costmap
    cost = make_costmap(4096, 4096, seed=1)
    h = capi.Handle(0); h.costmap_upload(cost); h.footprint_from_boundary(0.05, [0.64, 0.46], True)
    out = {}
    for nb in (72, 18, 12, 9):
        for _ in range(3): h.se2_volume_build(72, 0.5, 0.5, 0.0, 0, nb)
        h.sync(); ts = []
        for _ in range(5):
            h.l2_flush(); h.timer_start(); h.se2_volume_build(72, 0.5, 0.5, 0.0, 0, nb); ts.append(h.timer_stop())
        out[nb] = (round(float(np.mean(ts)), 3), round(h.se2_volume_stats()["samples_per_cell"], 1))
    print(sys.argv[1], out)
else:
    for n in (2, 3, 4, 5, 6, 8):
        env = dict(os.environ, PP_SE2_MAX_TABLES=str(n), PP_SE2_FEW_TABLES=str(n))
        subprocess.run([sys.executable, __file__, str(n)], env=env)
