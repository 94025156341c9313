"""CPU: the oracle restatement (oracle/port.c) against the golden vectors produced by the
reference's own code, and against that code itself where oracle/_ref is built."""
import numpy as np
import pytest

from conftest import FOOTPRINTS, OFFSETS, RES
from path_planner_b200.synth import make_costmap, make_teb_scene


@pytest.mark.parametrize("fname", list(FOOTPRINTS))
def test_footprint_points_match_golden(port, golden, fname):
    b, hole = FOOTPRINTS[fname]
    assert np.array_equal(port.footprint_points(RES, b, hole), golden[f"fp_{fname}"])


def test_footprint_point_counts(port):
    # 15x11 box rim = 48, solid = 165 at 0.05 m; 28 / 63 at 0.1 m
    assert len(port.footprint_points(0.05, [0.64, 0.46], True)) == 48
    assert len(port.footprint_points(0.05, [0.64, 0.46], False)) == 165
    assert len(port.footprint_points(0.1, [0.64, 0.46], True)) == 28
    assert len(port.footprint_points(0.1, [0.64, 0.46], False)) == 63


@pytest.mark.parametrize("fname,oname", [(f, o) for f in FOOTPRINTS for o in OFFSETS if f == "rect_hole" or o == "centre"])
def test_volume_matches_golden(port, golden, fname, oname):
    b, hole = FOOTPRINTS[fname]
    pts = port.footprint_points(RES, b, hole)
    v = port.se2_volume(golden["cost"], pts, 72, *OFFSETS[oname])
    assert np.array_equal(v, golden[f"vol_{fname}_{oname}_72"])


def test_volume_144_bins_matches_golden(port, golden):
    pts = port.footprint_points(RES, [0.64, 0.46], True)
    bins = golden["vol_rect_hole_centre_144_bins"]
    for k, t in enumerate(bins):
        v = port.se2_volume(golden["cost"], pts, 144, 0.5, 0.5, 0.0, t0=int(t), t1=int(t) + 1)
        assert np.array_equal(v[0], golden["vol_rect_hole_centre_144"][k])


def test_batch_matches_golden(port, golden):
    pts = port.footprint_points(RES, [0.64, 0.46], True)
    c, f, k = port.footprint_cost_batch(golden["cost"], pts, golden["batch_poses"])
    assert np.array_equal(c, golden["batch_cost"])
    assert np.array_equal(f, golden["batch_flag"])
    assert np.array_equal(k, golden["batch_collision"])
    assert np.array_equal(f, k)  # checkFootPrintCollision == flag of checkAndGetFootprintCost


def test_heuristic_matches_golden(port, golden):
    ok, table, qo = port.heuristic(golden["cost"], golden["heur_start"], golden["heur_goal"], 2.0, golden["heur_queries"])
    assert int(ok) == int(golden["heur_ok"])
    assert np.array_equal(table, golden["heur_table"])
    assert np.array_equal(qo, golden["heur_query_out"])
    ok2, table2, _ = port.heuristic(golden["cost"], (0.0, 0.0), golden["heur_goal"], 2.0)
    assert int(ok2) == int(golden["heur_full_ok"]) == 0
    assert np.array_equal(table2, golden["heur_full_table"])


def test_fixed_point_vs_reference_table(port, golden):
    """The converged relaxation equals the reference's closed values up to the 1-ulp effect of its
    A* tie-breaking; never above, and far inside the 1e-4 gate."""
    g = golden["heur_goal"]
    fp = port.heuristic_fixed_point(golden["cost"], (int(g[0]), int(g[1])), 2.0)
    ref_table = golden["heur_full_table"]
    closed = ref_table > 0
    assert np.array_equal(closed, fp > 0)
    assert (ref_table[closed] >= fp[closed]).all()
    rel = np.abs(ref_table[closed] - fp[closed]) / fp[closed]
    assert rel.max() <= 2.4e-7
    assert (fp[~closed] == 0).all()
    assert fp[int(g[1]), int(g[0])] == np.float32(0.00001)


def test_radius_topk_matches_golden(port, golden):
    idx, d2, cnt, found = port.radius_topk(golden["teb_pts"], golden["teb_queries"], 0.9574, 10)
    assert np.array_equal(cnt, golden["teb_cnt"])
    assert np.array_equal(found, golden["teb_found"])
    assert np.array_equal(d2, golden["teb_d2"])
    # equal distances may be ordered differently by nanoflann's unstable sort
    same = idx == golden["teb_idx"]
    if not same.all():
        q, j = np.nonzero(~same)
        for a, b in zip(q, j):
            assert (golden["teb_d2"][a] == d2[a, b]).sum() > 1


def test_extract_obstacles_order_and_filters(port):
    cost = make_costmap(60, 50, seed=2)
    pts = port.extract_obstacles(cost, 0.05, (-1.0, -0.5), (0.4, 0.6, 0.3), 4.5, 1.5)
    assert len(pts) > 0
    # x-outer / y-inner order (teb_follower.cpp:148-150): x non-decreasing, y increasing within a column
    assert (np.diff(pts[:, 0]) >= 0).all()
    same_col = np.diff(pts[:, 0]) == 0
    assert (np.diff(pts[:, 1])[same_col] > 0).all()
    # no candidate further behind than the ignore range
    d = pts - np.array([0.4, 0.6])
    behind = d @ np.array([np.cos(0.3), np.sin(0.3)]) < 0
    assert (np.hypot(d[:, 0], d[:, 1])[behind] <= 1.5).all()


# ---- against the reference's own code, where it is built (this container) ---------------------

def test_port_equals_reference_volume(port, ref):
    if ref is None:
        pytest.skip("oracle/_ref not built on this machine")
    cost = make_costmap(130, 70, seed=21)
    for fname, (b, hole) in FOOTPRINTS.items():
        pts = port.footprint_points(RES, b, hole)
        assert np.array_equal(pts, ref.footprint_points(RES, b, hole))
        for off in OFFSETS.values():
            a = port.se2_volume(cost, pts, 72, *off, t0=0, t1=72)
            r = ref.se2_volume(cost, RES, b, hole, 72, *off)
            assert np.array_equal(a, r)


def test_port_equals_reference_heuristic_and_radius(port, ref):
    if ref is None:
        pytest.skip("oracle/_ref not built on this machine")
    cost = make_costmap(150, 120, seed=4, coverage=0.08, pockets=[(12, 12, 5), (140, 110, 5)])
    rng = np.random.default_rng(1)
    q = np.stack([rng.uniform(1, 149, 30), rng.uniform(1, 119, 30)], 1).astype(np.float32)
    okr, tr, qr = ref.heuristic(cost, RES, (12.5, 12.5), (140.5, 110.5), 2.1, q)
    okp, tp, qp = port.heuristic(cost, (12.5, 12.5), (140.5, 110.5), 2.1, q)
    assert okr == okp and np.array_equal(tr, tp) and np.array_equal(qr, qp)
    pts, poses = make_teb_scene(3000, 50, seed=9)
    r = ref.radius_topk(pts, poses, 0.9574, 5)
    p = port.radius_topk(pts, poses, 0.9574, 5)
    assert np.array_equal(r[2], p[2]) and np.array_equal(r[3], p[3]) and np.array_equal(r[1], p[1])


def test_double_trig_variant_differs_rarely(port):
    """The alternative overload resolution (double cos/sin) is a different
    function: it agrees almost everywhere but not bit-for-bit, which is why the switch exists."""
    cost = make_costmap(200, 120, seed=8)
    pts = port.footprint_points(RES, [0.64, 0.46], True)
    a = port.se2_volume(cost, pts, 72, 0.0, 0.0, 0.0)
    b = port.se2_volume(cost, pts, 72, 0.0, 0.0, 0.0, trig_double=True)
    frac = (a != b).mean()
    assert frac < 0.02
