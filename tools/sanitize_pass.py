"""One small-map pass through every kernel of the library, for compute-sanitizer (profiles/sanitizer_r2.md):
   compute-sanitizer --tool memcheck|racecheck|initcheck python tools/sanitize_pass.py
Checks results against the CPU checker as it goes, so a sanitizer run is also a (slow) parity run."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from path_planner_b200 import capi
from path_planner_b200.synth import make_costmap, make_teb_scene

port = oracle.load_port()
W, H = 288, 80   # three 128-wide tiles (the last one partial) x five 16-row tiles
cost = make_costmap(W, H, seed=4, coverage=0.1, pockets=[(30, 30, 4)])
h = capi.Handle(0)
h.costmap_upload(cost)
# K1: every compiled halo of the tiled kernel that a footprint of this size reaches (guard-free tables, mixed
# tiles at the first columns / rows), bulk-tensor stores and plain stores, then the generic kernel
for name, (b, hole) in {"rect": ([0.64, 0.46], True), "box": ([-0.2, -0.25, 0.6, 0.25], True), "circle": ([0.35], True),
                        "solid": ([0.64, 0.46], False), "big": ([1.2, 0.9], True), "small": ([0.2, 0.2], True)}.items():
    h.footprint_from_boundary(0.05, b, hole)
    pts = port.footprint_points(0.05, b, hole)
    want = port.se2_volume(cost, pts, 72)
    for mode in (0, 1):
        h.se2_volume_set_mode(mode)
        h.se2_volume_build(72)
        assert np.array_equal(h.se2_volume_download(), want), (name, mode)
    h.se2_volume_set_mode(0)
    os.environ["PP_SE2_NO_TMA"] = "1"
    h.se2_volume_build(72)
    assert np.array_equal(h.se2_volume_download(), want), (name, "plain stores")
    del os.environ["PP_SE2_NO_TMA"]
    h.se2_volume_build_bins(72, [3, 39, 20, 56, 71])
    print("K1", name, "ok", h.se2_volume_stats())
# dirty-region refresh
h.footprint_from_boundary(0.05, [0.64, 0.46], True)
h.se2_volume_build(72)
c2 = cost.copy()
c2[20:40, 100:140] = 254
h.costmap_update_rect(100, 20, np.ascontiguousarray(c2[20:40, 100:140]))
h.se2_volume_refresh()
assert np.array_equal(h.se2_volume_download(), port.se2_volume(c2, port.footprint_points(0.05, [0.64, 0.46], True), 72))
h.costmap_upload(cost)
# K1b batch
poses = np.random.default_rng(0).uniform([-5, -5, -4], [W + 5, H + 5, 4], (500, 3)).astype(np.float32)
c, f = h.footprint_cost_batch(poses)
oc, of, _ = port.footprint_cost_batch(cost, port.footprint_points(0.05, [0.64, 0.46], True), poses)
assert np.array_equal(c, oc) and np.array_equal(f, of)
# K2: with programmatic stream serialization (the handle's entry point) and without (the no-alloc entry point)
field = h.heuristic_build(30, 30, 2.0)
assert np.array_equal(field, port.heuristic_fixed_point(cost, (30, 30), 2.0))
import torch
d_cost = torch.from_numpy(cost).cuda()
d_field = torch.empty((H, W), dtype=torch.float32, device="cuda")
capi.heuristic_build_noalloc(d_cost.data_ptr(), W, H, 30, 30, 2.0, d_field.data_ptr(), torch.cuda.current_stream().cuda_stream)
assert np.array_equal(d_field.cpu().numpy(), field)
# K3
ext = h.obstacles_extract(0.05, (0.0, 0.0), (5.0, 2.0, 0.3), 4.5, 1.5)
q = np.stack([np.linspace(1, 12, 40), np.linspace(0.5, 3.5, 40)], 1)
got = h.radius_topk(None, q, 0.9574, 10)
o = port.radius_topk(ext, q, 0.9574, 10)
assert np.array_equal(got[1], o[1]) and np.array_equal(got[2], o[2])
pts, poses2 = make_teb_scene(6000, 20, seed=1)
dense = np.concatenate([pts, np.random.default_rng(1).normal(0, 0.3, (4000, 2))])
got = h.radius_topk(dense, poses2, 0.9574, 10)
o = port.radius_topk(dense, poses2, 0.9574, 10)
assert np.array_equal(got[1], o[1]) and np.array_equal(got[2], o[2])
# N2 ingestion, N3 edges
occ = np.where(cost == 254, 100, 0).astype(np.int8)
h.occupancy_upload(occ)
h.costmap_from_occupancy(capi.LayerParams.make())
assert np.array_equal(h.costmap_download(), port.costmap_from_occupancy(occ, oracle.layer_params(), "nearest"))
poses3 = np.concatenate([poses2, np.linspace(-3.0, 3.0, len(poses2))[:, None]], axis=1)
h.radius_topk(pts, poses2, 0.9574, 10)
h.robot_model_set([0.72, 0.50])
err, jac = h.obstacle_edges(poses3, 10)
# N1 plan (host search over the tables; volume paging is off at this size)
hp = capi.Handle(0)
pc = make_costmap(400, 400, seed=3, coverage=0.06, unknown_frac=0.0, pockets=[(40, 40, 25), (350, 340, 25)])
hp.costmap_upload(pc)
hp.footprint_from_boundary(0.05, [0.64, 0.46], True)
r = hp.plan(0.05, (0.0, 0.0), (2.0, 2.0, 0.3), (17.5, 17.0, 1.2), {"motion_type": 2})
print("plan found", r["ok"], r["iterations"])
hp.close()
h.close()
print("sanitize pass ok")
