"""dev probe: where does the fast volume differ from the oracle port (small map)?"""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1:
    import numpy as np
    import oracle
    from path_planner_b200 import capi
    from path_planner_b200.synth import make_costmap
    port = oracle.load_port()
    W, H = int(sys.argv[2]), int(sys.argv[3])
    cost = make_costmap(W, H, seed=3, coverage=0.1)
    h = capi.Handle(0); h.costmap_upload(cost); h.footprint_from_boundary(0.05, [0.64, 0.46], True)
    h.se2_volume_build(72)
    v = h.se2_volume_download()
    want = port.se2_volume(cost, port.footprint_points(0.05, [0.64, 0.46], True), 72)
    bad = np.argwhere(v != want)
    print("tables", sys.argv[1], "map", W, H, "mismatches", len(bad), "of", v.size, h.se2_volume_stats())
    if len(bad):
        print(" bins", np.unique(bad[:, 0])[:20], " x%32 hist", np.bincount(bad[:, 2] % 32, minlength=32).tolist())
        print(" x/32 hist", np.bincount(bad[:, 2] // 32).tolist(), " y%16 hist", np.bincount(bad[:, 1] % 16, minlength=16).tolist())
        for b in bad[:8]:
            print("  ", b.tolist(), "got", v[tuple(b)], "want", want[tuple(b)])
else:
    for n in ("1", "2", "3", "12"):
        for (W, H) in ((128, 16), (256, 64), (96, 80)):
            subprocess.run([sys.executable, __file__, n, str(W), str(H)], env=dict(os.environ, PP_SE2_MAX_TABLES=n))
