"""GPU: the north star's third correctness gate -- with the GPU tables substituted in, the
reference's own Hybrid-A* search (astar_hybrid.cpp, compiled from the reference sources with one
hook line, see oracle/e2e_harness.cpp) returns the identical path.
  (A) oracle-built volume + the reference's own lazily grown heuristic
  (B) GPU-built volume + GPU heuristic field injected
Pass = identical Coordinates and identical iteration count. The unmodified reference (C) is run
for context: it checks continuous poses, (A)/(B) quantise them to cell centre and nearest bin."""
import numpy as np
import pytest

import oracle
from path_planner_b200.synth import make_costmap

pytestmark = pytest.mark.gpu
RES = 0.05


def scene(W, seed, coverage, start, goal, radius=25):
    """Planning scenes: no speckled NO_INFORMATION cells (any 255 under the footprint is a collision,
    collision_checker.cpp:165) and start/goal pockets wider than the footprint reach."""
    pockets = [(int(start[0] / RES), int(start[1] / RES), radius), (int(goal[0] / RES), int(goal[1] / RES), radius)]
    return make_costmap(W, W, seed=seed, coverage=coverage, unknown_frac=0.0, pockets=pockets)


def run_case(cost, boundary, hole, start, goal, params, port):
    e2e = oracle.load_e2e()
    if e2e is None:
        pytest.skip("oracle/_ref/libpp_e2e.so not built (needs the reference sources at build time)")
    n_theta = params.get("n_theta", 72)
    frac = e2e.heading_fraction(start[2], n_theta)
    pts = port.footprint_points(RES, boundary, hole)
    vol = port.se2_volume(cost, pts, n_theta, 0.5, 0.5, frac, nthreads=0)
    a = e2e.plan(0, cost, RES, boundary, hole, start, goal, oracle_volume=vol, params=params)
    b = e2e.plan(1, cost, RES, boundary, hole, start, goal, params=params)
    return a, b


def check_identical(a, b):
    assert a["ok"] == 1 and b["ok"] == 1, (a["ok"], b["ok"])
    assert a["iterations"] == b["iterations"], (a["iterations"], b["iterations"])
    assert a["path"].shape == b["path"].shape
    assert np.array_equal(a["path"], b["path"])
    assert b["exact_calls"] == 0 and b["lookups"] > 0  # every footprint check was a table lookup


def test_identical_path_config1_400x400_rect(port):
    """BASELINE config 1: 400x400 @0.05 m, 72 bins, rectangular footprint, Reeds-Shepp."""
    start, goal = (2.0, 2.0, 0.3), (17.5, 17.0, 1.2)
    a, b = run_case(scene(400, 3, 0.06, start, goal), [0.64, 0.46], True, start, goal, {"motion_type": 2}, port)
    check_identical(a, b)
    assert len(a["path"]) > 50 and a["iterations"] > 10000  # a long search: ~9e4 expansions
    print("config1: iterations %d, plan %.1f ms (A) vs %.1f ms (B; prepare %.1f ms)" % (a["iterations"], a["plan_ms"], b["plan_ms"], b["prepare_ms"]))


@pytest.mark.parametrize("seed,start,goal", [(5, (1.5, 1.5, 0.0), (18.0, 3.0, -1.0)), (8, (3.0, 17.0, -0.5), (16.0, 2.5, 0.0)),
                                             (9, (10.0, 10.0, 2.0), (2.0, 18.0, 3.0))])
def test_identical_path_more_scenes(port, seed, start, goal):
    a, b = run_case(scene(400, seed, 0.05, start, goal), [0.64, 0.46], True, start, goal, {"motion_type": 2}, port)
    if a["ok"] != 1:
        assert b["ok"] == a["ok"] and a["iterations"] == b["iterations"]  # both fail the same way
        return
    check_identical(a, b)


def test_identical_path_config2_2048_box(port):
    """BASELINE config 2: 2048x2048 with random obstacles, 72 bins, the asymmetric 4-number box
    (the reference checker's closest thing to a polygon footprint)."""
    start, goal = (5.0, 5.0, 0.5), (60.0, 45.0, 0.0)
    cost = scene(2048, 2, 0.04, start, goal)
    a, b = run_case(cost, [-0.2, -0.25, 0.6, 0.25], True, start, goal, {"motion_type": 2, "max_iterations": 400000}, port)
    check_identical(a, b)
    print("config2: iterations %d, plan %.1f ms (A, CPU tables) vs %.1f ms (B, GPU tables; prepare %.1f ms)"
          % (a["iterations"], a["plan_ms"], b["plan_ms"], b["prepare_ms"]))


def test_unmodified_reference_for_context(port, ref):
    """(C): not a parity gate -- shows how far the cell-quantised tables move the plan."""
    if ref is None:
        pytest.skip("oracle/_ref not built")
    start, goal = (1.5, 1.5, 0.0), (18.0, 3.0, -1.0)
    cost = scene(400, 5, 0.05, start, goal)
    ok, path, it, ms = ref.plan(cost, RES, [0.64, 0.46], True, start, goal, {"motion_type": 2})
    a, b = run_case(cost, [0.64, 0.46], True, start, goal, {"motion_type": 2}, port)
    assert ok and b["ok"] == 1
    assert np.allclose(path[0], b["path"][0]) and np.allclose(path[-1], b["path"][-1])  # same endpoints
    print("reference: %d iterations %.1f ms; table-driven: %d iterations %.1f ms (+%.1f ms prepare)"
          % (it, ms, b["iterations"], b["plan_ms"], b["prepare_ms"]))


@pytest.mark.parametrize("W,seed,cov,start,goal,boundary,extra", [
    (400, 3, 0.06, (2.0, 2.0, 0.3), (17.5, 17.0, 1.2), [0.64, 0.46], {}),
    (400, 5, 0.05, (1.5, 1.5, 0.0), (18.0, 3.0, -1.0), [0.64, 0.46], {}),
    (400, 8, 0.05, (3.0, 17.0, -0.5), (16.0, 2.5, 0.0), [0.64, 0.46], {}),
    (400, 9, 0.05, (10.0, 10.0, 2.0), (2.0, 18.0, 3.0), [0.64, 0.46], {"motion_type": 1}),
    (2048, 2, 0.04, (5.0, 5.0, 0.5), (60.0, 45.0, 0.0), [-0.2, -0.25, 0.6, 0.25], {"max_iterations": 400000}),
])
def test_product_search_matches_reference_search(handle, port, W, seed, cov, start, goal, boundary, extra):
    """pp_plan (the library's own host Hybrid-A* over the GPU tables, host/hybrid_astar.cpp) against
    the reference's search code running over the same kind of tables: same path, same iterations."""
    e2e = oracle.load_e2e()
    if e2e is None:
        pytest.skip("oracle/_ref/libpp_e2e.so not built")
    params = {"motion_type": 2}
    params.update(extra)
    cost = scene(W, seed, cov, start, goal)
    ref_run = e2e.plan(1, cost, RES, boundary, True, start, goal, params=params)
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, boundary, True)
    got = handle.plan(RES, (0.0, 0.0), start, goal, params)
    assert got["ok"] == ref_run["ok"]
    assert got["iterations"] == ref_run["iterations"]
    assert got["path"].shape == ref_run["path"].shape
    assert np.array_equal(got["path"], ref_run["path"])
    print("pp_plan %dx%d: %d iterations, tables %.1f ms + search %.1f ms (reference search on tables: %.1f ms)"
          % (W, W, got["iterations"], got["tables_ms"], got["search_ms"], ref_run["plan_ms"]))


def test_plan_rejects_and_fails_cleanly(handle):
    from path_planner_b200.capi import PPError
    with pytest.raises(PPError):
        handle.plan(RES, (0.0, 0.0), (1.0, 1.0, 0.0), (2.0, 2.0, 0.0))  # no costmap
    cost = np.zeros((100, 100), np.uint8)
    cost[:, 50] = 254  # a wall: the goal side is unreachable
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, [0.2], False)
    r = handle.plan(RES, (0.0, 0.0), (1.0, 2.5, 0.0), (4.0, 2.5, 0.0))
    assert r["ok"] == 0 and len(r["path"]) == 0
    r = handle.plan(RES, (0.0, 0.0), (1.0, 2.5, 0.0), (40.0, 2.5, 0.0))  # goal outside the map
    assert r["ok"] == 0
    r = handle.plan(RES, (0.0, 0.0), (0.6, 0.6, 0.0), (1.9, 1.2, 0.5))
    assert r["ok"] == 1 and len(r["path"]) > 2


EXACT_SCENES = [
    (400, 3, 0.06, (2.0, 2.0, 0.3), (17.5, 17.0, 1.2), [0.64, 0.46], {}),
    (400, 5, 0.05, (1.5, 1.5, 0.0), (18.0, 3.0, -1.0), [0.64, 0.46], {}),
    (400, 9, 0.05, (10.0, 10.0, 2.0), (2.0, 18.0, 3.0), [0.64, 0.46], {"motion_type": 1}),
    (2048, 2, 0.04, (5.0, 5.0, 0.5), (60.0, 45.0, 0.0), [-0.2, -0.25, 0.6, 0.25], {"max_iterations": 400000}),
]


@pytest.mark.parametrize("W,seed,cov,start,goal,boundary,extra", EXACT_SCENES)
def test_exact_analytic_expansion_matches_reference_search(handle, W, seed, cov, start, goal, boundary, extra):
    """PP_PLAN_EXACT_ANALYTIC: the samples of analytic expansions are decided at their continuous poses, as the
    reference's checkFootPrintCollision does (collision_checker.cpp:126-142). Checker: the reference's own search
    code over the drop-in CollisionChecker with setExactAnalytic(true), which asks pp_footprint_cost_batch once
    per sample. The product filters with the volume, re-checks flagged samples on the host and confirms a
    surviving expansion with one batched launch: same path, same iterations."""
    e2e = oracle.load_e2e()
    if e2e is None:
        pytest.skip("oracle/_ref/libpp_e2e.so not built")
    params = {"motion_type": 2}
    params.update(extra)
    cost = scene(W, seed, cov, start, goal)
    ref_run = e2e.plan(1 + 4, cost, RES, boundary, True, start, goal, params=params)
    assert ref_run["exact_calls"] > 0
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, boundary, True)
    got = handle.plan(RES, (0.0, 0.0), start, goal, dict(params, exact_analytic=True, validate_path=True))
    assert got["ok"] == ref_run["ok"]
    assert got["iterations"] == ref_run["iterations"]
    assert np.array_equal(got["path"], ref_run["path"])
    st = got["stats"]
    # far fewer exact evaluations than one launch per sample: a handful of batches per plan
    assert st["analytic_confirm_batches"] <= max(4, st["analytic_attempts"] // 10), st
    print("exact analytic %dx%d: %d iterations, %s; reference-side per-sample exact calls: %d" % (W, W, got["iterations"], st, ref_run["exact_calls"]))


@pytest.mark.parametrize("W,seed,cov,start,goal,boundary,extra", EXACT_SCENES)
def test_quantisation_of_the_volume_lookup_is_bounded(handle, W, seed, cov, start, goal, boundary, extra):
    """Every footprint question of a search (continuous pose), answered exactly in ONE pp_footprint_cost_batch
    call and compared with what the volume said (cell centre, nearest bin). The rates are reported in DESIGN.md;
    the bounds here pin them. The product's log must also equal the reference search's own questions."""
    params = {"motion_type": 2}
    params.update(extra)
    cost = scene(W, seed, cov, start, goal)
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, boundary, True)
    handle.plan_set_query_log(True)
    got = handle.plan(RES, (0.0, 0.0), start, goal, dict(params, validate_path=True))
    poses, kind, qflag, qcost = handle.plan_query_log()
    handle.plan_set_query_log(False)
    assert len(poses) == got["stats"]["lookups"] and len(poses) > 100
    e2e = oracle.load_e2e()
    if e2e is not None:
        e2e.plan(1 + 8, cost, RES, boundary, True, start, goal, params=params)
        rp, rk, rf, rc = e2e.query_log()
        assert np.array_equal(rp, poses) and np.array_equal(rk, kind) and np.array_equal(rf, qflag) and np.array_equal(rc, qcost)
    xcost, xflag = handle.footprint_cost_batch(poses)
    for k, name in ((0, "primitives"), (1, "analytic samples")):
        m = kind == k
        if not m.any():
            continue
        false_free = float(((qflag == 0) & (xflag == 1) & m).sum()) / m.sum()
        false_coll = float(((qflag == 1) & (xflag == 0) & m).sum()) / m.sum()
        cost_differs = float(((qcost != xcost.astype(np.uint8)) & m).sum()) / m.sum()
        print("quantisation %dx%d %s: %d questions, false-free %.4f, false-collision %.4f, cost differs %.3f"
              % (W, W, name, m.sum(), false_free, false_coll, cost_differs))
        assert false_free < 0.03 and false_coll < 0.03
    print("path poses colliding at their continuous pose: %d of %d" % (got["stats"]["path_poses_in_collision"], got["stats"]["path_poses_checked"]))
    assert got["stats"]["path_poses_checked"] == len(got["path"])


def test_path_pose_that_fails_the_exact_check_is_planned_around(handle, monkeypatch):
    """PP_PLAN_VALIDATE_PATH: a pose of the found path that collides at its continuous position has its search
    node excluded and the search is repeated. No test scene produces such a pose (0 of 163 .. 937), so the
    mechanism is driven by a debug knob that declares pose 40 of the first path colliding."""
    start, goal = (1.5, 1.5, 0.0), (18.0, 3.0, -1.0)
    cost = scene(400, 5, 0.05, start, goal)
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, [0.64, 0.46], True)
    params = {"motion_type": 2, "validate_path": True}
    first = handle.plan(RES, (0.0, 0.0), start, goal, params)
    assert first["ok"] == 1 and first["stats"]["replans"] == 0 and first["stats"]["path_poses_in_collision"] == 0
    monkeypatch.setenv("PP_DEBUG_PLAN_FAKE_HIT", "40")
    second = handle.plan(RES, (0.0, 0.0), start, goal, params)
    monkeypatch.delenv("PP_DEBUG_PLAN_FAKE_HIT")
    assert second["ok"] == 1 and second["stats"]["replans"] == 1
    assert second["stats"]["path_poses_in_collision"] == 0          # the second path passes the exact check
    assert not any(np.array_equal(first["path"][40], p) for p in second["path"])  # and avoids the excluded pose
    assert np.array_equal(second["path"][0], first["path"][0]) and np.array_equal(second["path"][-1], first["path"][-1])
