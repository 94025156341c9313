"""dev probe: per-visit timeline of one heuristic build (needs the -DPP_HEUR_TIMELINE build:
   make -C path_planner_b200/csrc BUILD=build_tl OUT=../libpp_b200_tl.so EXTRA="-DPP_SE2_TIMELINE -DPP_HEUR_TIMELINE").
   PP_B200_LIB=path_planner_b200/libpp_b200_tl.so python tools/k2_timeline.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from path_planner_b200 import capi
from path_planner_b200.synth import make_costmap
cost = make_costmap(4096, 4096, seed=1, pockets=[(3000, 1000, 6)])
h = capi.Handle(0); h.costmap_upload(cost)
path = "/tmp/pp_heur_tl.txt"
os.environ["PP_HEUR_TIMELINE_FILE"] = path  # every build writes its own visits and resets the log
for _ in range(2):
    h.heuristic_build(3000, 1000, 2.0, download=False)
h.sync()
h.timer_start(); h.heuristic_build(3000, 1000, 2.0, download=False); ms = h.timer_stop()
tl = np.loadtxt(path, dtype=np.uint64).astype(np.int64)
print(f"build {ms:.3f} ms, {len(tl)} visits recorded", h.heuristic_stats())
rnd = tl[:, 0]
t0 = tl[:, 2].min()
T = (tl[:, 2:8] - t0) / 1e3  # entry, wait, list, staged, converged, published
rows = []
for r in np.unique(rnd):
    m = rnd == r
    e = T[m]
    rows.append((r, m.sum(), e[:, 0].min(), e[:, 1].min(), e[:, 1].max(), e[:, 2].max(), e[:, 3].max(), e[:, 4].max(), e[:, 5].max(), tl[m, 8].max(), tl[m, 8].mean()))
rows = np.array(rows)
# per round: span from the previous round's last publish to this round's last publish
end = rows[:, 8]
span = np.diff(end)
print("rounds", len(rows), " median round span %.1f us, mean %.1f us" % (np.median(span), span.mean()))
def stat(name, v):
    print(f"  {name:46s} median {np.median(v):6.2f}  mean {v.mean():6.2f}  p90 {np.percentile(v, 90):6.2f} us")
prev_end = end[:-1]
cur = rows[1:]
stat("prev round's last publish -> first wait done", cur[:, 3] - prev_end)
stat("first wait done -> last wait done", cur[:, 4] - cur[:, 3])
stat("last wait done -> last list read", cur[:, 5] - cur[:, 4])
stat("last list read -> last staged", cur[:, 6] - cur[:, 5])
stat("last staged -> last converged", cur[:, 7] - cur[:, 6])
stat("last converged -> last published", cur[:, 8] - cur[:, 7])
# per visit phases
stat("per visit: wait done -> list read", T[:, 2] - T[:, 1])
stat("per visit: staging (incl. earlier visits of the CTA)", T[:, 3] - T[:, 2])
stat("per visit: sweeps", T[:, 4] - T[:, 3])
stat("per visit: publish", T[:, 5] - T[:, 4])
print("  sweeps (votes) per visit: median %.0f mean %.1f max %d; per round max: median %.0f" % (np.median(tl[:, 8]), tl[:, 8].mean(), tl[:, 8].max(), np.median(rows[:, 9])))
print("  time per vote in the slowest visit of a round: median %.2f us" % np.median((cur[:, 7] - cur[:, 6]) / np.maximum(1, cur[:, 9])))
print("  tiles per round: median %d max %d" % (np.median(rows[:, 1]), rows[:, 1].max()))
for r in rows[::15]:
    print("   round %3d tiles %4d: entry %.1f wait %.1f..%.1f list %.1f staged %.1f conv %.1f pub %.1f  votes max %d mean %.1f" % tuple(r))
