"""GPU parity: K2 heuristic field. Gate from the north star: within 1e-4 relative of the reference's
cached-A* costs. Measured: bit-equal to the converged relaxation (oracle fixed point), hence
<= 1 ulp below the reference's closed values."""
import numpy as np
import pytest

from path_planner_b200.synth import make_costmap

pytestmark = pytest.mark.gpu
REL_TOL = 1e-4  # north-star tolerance


def check_against_reference_table(field, table):
    closed = table > 0
    assert np.array_equal(field > 0, closed)
    rel = np.abs(field[closed] - table[closed]) / table[closed]
    assert rel.max() <= REL_TOL
    assert rel.max() <= 2.4e-7  # what is actually observed: at most 1 ulp
    assert (field[~closed] == 0).all()


def test_field_matches_golden_reference_tables(handle, golden):
    handle.costmap_upload(golden["cost"])
    g = golden["heur_goal"]
    field = handle.heuristic_build(int(g[0]), int(g[1]), 2.0)
    check_against_reference_table(field, golden["heur_full_table"])
    # the lazily grown table of a real plan closes only part of the map; where it did, same values
    lazy = golden["heur_table"]
    closed = lazy > 0
    rel = np.abs(field[closed] - lazy[closed]) / lazy[closed]
    assert rel.max() <= 2.4e-7


@pytest.mark.parametrize("shape,seed", [((96, 80), 1), ((131, 77), 2), ((700, 500), 3), ((33, 1025), 4), ((1024, 64), 5)])
def test_field_bit_equal_to_fixed_point(handle, port, shape, seed):
    W, H = shape
    gx, gy = W // 3, H // 2
    cost = make_costmap(W, H, seed=seed, coverage=0.1, pockets=[(gx, gy, 3)])
    handle.costmap_upload(cost)
    for pen in (2.0, 2.1, 0.0):
        field = handle.heuristic_build(gx, gy, pen)
        fp = port.heuristic_fixed_point(cost, (gx, gy), pen)
        assert np.array_equal(field, fp), (shape, pen, handle.heuristic_stats())


def test_field_vs_reference_astar_on_larger_map(handle, port):
    cost = make_costmap(600, 400, seed=7, coverage=0.08, pockets=[(500, 350, 4)])
    _, table, _ = port.heuristic(cost, (0.0, 0.0), (500.5, 350.5), 2.0)  # bit-equal to the reference's code
    handle.costmap_upload(cost)
    field = handle.heuristic_build(500, 350, 2.0)
    check_against_reference_table(field, table)


def test_unreachable_and_degenerate_goals(handle, port):
    cost = np.zeros((64, 64), np.uint8)
    cost[:, 30] = 254  # wall: right half unreachable from the left
    handle.costmap_upload(cost)
    field = handle.heuristic_build(5, 5, 2.0)
    assert (field[:, 31:] == 0).all() and (field[1:-1, 1:30] > 0).all()
    assert np.array_equal(field, port.heuristic_fixed_point(cost, (5, 5), 2.0))
    # goal inside an obstacle and on the border still seeds the field (astar_hybrid.cpp:116-117)
    for g in [(30, 10), (0, 0), (63, 20)]:
        field = handle.heuristic_build(g[0], g[1], 2.0)
        assert np.array_equal(field, port.heuristic_fixed_point(cost, g, 2.0)), g
    assert field[20, 63] == np.float32(0.00001)


def test_repeated_builds_with_goals_in_many_tiles(handle, port):
    """Every build must restart from a clean activity state, whichever tile holds the goal."""
    cost = make_costmap(520, 390, seed=13, coverage=0.06)
    handle.costmap_upload(cost)
    rng = np.random.default_rng(4)
    for _ in range(12):
        gx, gy = int(rng.integers(1, 519)), int(rng.integers(1, 389))
        field = handle.heuristic_build(gx, gy, 2.0)
        assert np.array_equal(field, port.heuristic_fixed_point(cost, (gx, gy), 2.0)), (gx, gy)
        if field.max() > 100.0:  # (a goal inside an obstacle reaches nothing: one round)
            assert handle.heuristic_stats()["rounds"] >= 8  # rounds that had work


def test_serpentine_corridor_and_cost_gradients(handle, port):
    """Maps on which the bands of a tile start and stop moving many times during one visit (the in-tile relaxation
    lets bands that have stopped sit votes out): a serpentine corridor of walls with 3-cell gaps, and a field of
    steep random costs without obstacles."""
    W, H = 257, 203
    cost = np.zeros((H, W), np.uint8)
    for i, x in enumerate(range(6, W - 6, 7)):
        cost[:, x] = 254
        gap = 2 if i % 2 == 0 else H - 5
        cost[gap:gap + 3, x] = 0
    handle.costmap_upload(cost)
    for goal in [(2, 2), (W - 3, H // 2), (130, 100)]:
        field = handle.heuristic_build(goal[0], goal[1], 2.0)
        assert np.array_equal(field, port.heuristic_fixed_point(cost, goal, 2.0)), goal
    rng = np.random.default_rng(9)
    steep = rng.integers(0, 253, size=(H, W)).astype(np.uint8)
    handle.costmap_upload(steep)
    for pen in (2.0, 7.5):
        field = handle.heuristic_build(100, 60, pen)
        assert np.array_equal(field, port.heuristic_fixed_point(steep, (100, 60), pen)), pen


def test_tiny_maps(handle, port):
    for shape in [(1, 1), (2, 5), (3, 3)]:
        cost = np.zeros(shape[::-1], np.uint8)
        handle.costmap_upload(cost)
        field = handle.heuristic_build(0, 0, 2.0)
        assert np.array_equal(field, port.heuristic_fixed_point(cost, (0, 0), 2.0))


def test_full_size_field_properties(handle):
    """4096x4096 (BASELINE config 3): properties that need no CPU run -- the recurrence holds at
    every reached cell (each value is exactly one rounded relaxation of its best neighbour) and
    values grow at least 1 per straight step from the goal."""
    W = H = 4096
    gx, gy = 3000, 1000
    cost = make_costmap(W, H, seed=1, pockets=[(gx, gy, 6)])
    handle.costmap_upload(cost)
    f = handle.heuristic_build(gx, gy, 2.0)
    assert f[gy, gx] == np.float32(0.00001)
    reached = f > 0
    assert reached.mean() > 0.5
    inf = np.float32(np.inf)
    F = np.where(reached, f, inf)
    Fp = np.pad(F, 1, constant_values=inf)
    s2 = np.sqrt(np.float32(2.0))
    scale = np.float32(1.0) + (np.float32(2.0) * cost.astype(np.float32) / np.float32(252.0))
    ms = np.minimum(np.minimum(Fp[:-2, 1:-1], Fp[2:, 1:-1]), np.minimum(Fp[1:-1, :-2], Fp[1:-1, 2:]))
    md = np.minimum(np.minimum(Fp[:-2, :-2], Fp[2:, 2:]), np.minimum(Fp[:-2, 2:], Fp[2:, :-2]))
    cand = np.minimum(ms + np.float32(1.0) * scale, md + s2 * scale)
    enterable = np.zeros((H, W), bool)
    enterable[1:-1, 1:-1] = cost[1:-1, 1:-1] < 253
    chk = reached & enterable
    chk[gy, gx] = False
    assert np.array_equal(f[chk], cand[chk])
    assert not (reached & ~enterable)[np.arange(H) != gy][:, np.arange(W) != gx].any()
    assert np.isinf(cand[enterable & ~reached]).all()  # an unreached cell has no reached neighbour
