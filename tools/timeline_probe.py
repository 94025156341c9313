"""dev probe: per-CTA timeline of one volume build (needs the -DPP_SE2_TIMELINE build:
   make -C path_planner_b200/csrc BUILD=build_tl OUT=../libpp_b200_tl.so EXTRA=-DPP_SE2_TIMELINE).
   PP_B200_LIB=path_planner_b200/libpp_b200_tl.so python tools/timeline_probe.py [rows]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 512
path = "/tmp/pp_timeline.txt"
from path_planner_b200 import capi
from path_planner_b200.synth import make_costmap
cost = make_costmap(4096, 4096, seed=1, pockets=[(3000, 1000, 6)])
h = capi.Handle(0); h.costmap_upload(cost); h.footprint_from_boundary(0.05, [0.64, 0.46], True)
y0 = 1024 if rows <= 2048 else 0
h.se2_volume_set_rows(y0, y0 + rows)
for _ in range(4):
    h.se2_volume_build(72)
h.sync()
h.timer_start(); e0 = h.timer_stop()
os.environ["PP_SE2_TIMELINE_FILE"] = path
h.l2_flush(); h.timer_start(); h.se2_volume_build(72); ms = h.timer_stop()
del os.environ["PP_SE2_TIMELINE_FILE"]
tl = np.loadtxt(path, dtype=np.uint64).astype(np.int64)
tl = tl[tl[:, 1] > 0]
t0 = tl[:, 1].min()
st, sg, bl, en, sm = (tl[:, 1] - t0) / 1e3, (tl[:, 2] - t0) / 1e3, (tl[:, 3] - t0) / 1e3, (tl[:, 4] - t0) / 1e3, tl[:, 5]
print(f"rows {rows}: {len(tl)} CTAs, event time {ms*1e3:.1f} us (empty event pair {e0*1e3:.1f} us); first start 0, last start {st.max():.1f}, last end {en.max():.1f} us")
order = np.argsort(st)
n = len(tl)
for lo in range(0, n, max(1, n // 12)):
    sel = order[lo:lo + max(1, n // 12)]
    print(f"  CTAs started {st[sel].min():7.1f}..{st[sel].max():7.1f} us: staging {np.mean(sg[sel]-st[sel]):5.1f}  build {np.mean(bl[sel]-sg[sel]):5.1f}  units {np.mean(en[sel]-bl[sel]):5.1f}  total {np.mean(en[sel]-st[sel]):5.1f} us")
# busy CTA slots over time
grid = np.arange(0, en.max(), 2.0)
act = [(np.sum((st <= t) & (en > t))) for t in grid]
print("  active CTAs every 2 us:", " ".join(str(a) for a in act))
print("  SMs used", len(np.unique(sm)), " end times per SM: min %.1f max %.1f" % (min(en[sm == s].max() for s in np.unique(sm)), en.max()))
late = np.argsort(-en)[:14]
print("  last to end: " + "; ".join(f"b{int(tl[i,0])} st {st[i]:.1f} sg {sg[i]-st[i]:.1f} bl {bl[i]-sg[i]:.1f} un {en[i]-bl[i]:.1f} end {en[i]:.1f}" for i in late))
tot = en - st
slow = np.argsort(-tot)[:14]
print(f"  median CTA {np.median(tot):.1f} us; longest: " + "; ".join(f"b{int(tl[i,0])} st {st[i]:.1f} sg {sg[i]-st[i]:.1f} bl {bl[i]-sg[i]:.1f} un {en[i]-bl[i]:.1f}" for i in slow))
