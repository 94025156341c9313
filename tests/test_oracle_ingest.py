"""CPU: the ingestion oracle (oracle/ingest.c) against its golden fixture, and the relation between
the two inflation definitions it holds (nearest-obstacle vs costmap_2d's queue propagation)."""
import os

import numpy as np
import pytest

from conftest import ROOT

import oracle


@pytest.fixture(scope="module")
def ingest_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "ingest_v1.npz"))


def random_occupancy(W, H, seed, p_lethal=0.01, p_unknown=0.05, gray=False):
    rng = np.random.default_rng(seed)
    occ = np.zeros((H, W), np.int8)
    if gray:
        occ[:] = rng.integers(0, 100, (H, W))
    occ[rng.random((H, W)) < p_lethal] = 100
    for _ in range(max(1, W * H // 20000)):  # some walls
        x, y, w, h = rng.integers(0, W), rng.integers(0, H), rng.integers(1, 40), rng.integers(1, 4)
        if rng.random() < 0.5:
            w, h = h, w
        occ[y:y + h, x:x + w] = 100
    occ[rng.random((H, W)) < p_unknown] = -1
    return occ


def test_threshold_matches_golden(port, ingest_golden):
    g = ingest_golden
    occ = port.occupancy_from_image(g["image"], float(g["occ_th"]), float(g["free_th"]), bool(g["negate"]))
    assert np.array_equal(occ, g["occ"])
    assert np.array_equal(port.occupancy_from_image(g["ramp"], float(g["occ_th"]), float(g["free_th"]), False),
                          g["ramp_occ"])
    # the reference compares the 0..255 pixel with the thresholds as given: sim.yaml's map is binary
    assert set(np.unique(g["occ"])) == {0, 100}
    assert set(np.unique(g["ramp_occ"])) > {0, 50, 100}


def test_layers_match_golden(port, ingest_golden):
    g = ingest_golden
    lp = oracle.layer_params(resolution=float(g["resolution"]))
    assert np.array_equal(port.costmap_from_occupancy(g["occ"], lp, "nearest"), g["cost_nearest"])
    assert np.array_equal(port.costmap_from_occupancy(g["occ"], lp, "propagate"), g["cost_propagate"])


def test_static_layer_values(port):
    occ = np.array([[-1, 0, 1, 50, 99, 100, 101, 127, -128, -2]], np.int8)
    tri = port.static_layer(occ, oracle.layer_params())
    assert tri.tolist() == [[255, 0, 0, 0, 0, 254, 254, 254, 254, 254]]
    assert port.static_layer(occ, oracle.layer_params(track_unknown_space=0)).tolist()[0][0] == 0
    scaled = port.static_layer(occ, oracle.layer_params(trinary_costmap=0))
    assert scaled.tolist() == [[255, 0, 2, 127, 251, 254, 254, 254, 254, 254]]


def test_cost_table_is_a_function_of_squared_distance(port):
    for lp in (oracle.layer_params(), oracle.layer_params(resolution=0.1), oracle.layer_params(resolution=0.025, inflation_radius=0.8),
               oracle.layer_params(cost_scaling_factor=3.0, inscribed_radius=0.3, inflation_radius=1.0)):
        R, tab = port.inflation_cost_table(lp)
        by_d2 = {}
        for i in range(R + 2):
            for j in range(R + 2):
                by_d2.setdefault(i * i + j * j, set()).add(int(tab[i, j]))
        assert all(len(v) == 1 for v in by_d2.values())
        d2 = sorted(by_d2)
        costs = [next(iter(by_d2[k])) for k in d2]
        assert costs == sorted(costs, reverse=True)  # never rises with distance
        assert tab[0, 0] == 254


def test_shadowed_obstacle_is_where_the_definitions_part(port):
    """Cell (20,20): nearest lethal cell at (+6,+6) (d^2 = 72); its two four-neighbours towards it belong
    to the obstacles at (+8,+3) and (+3,+8) (d^2 = 73 from the cell), so costmap_2d's queue never offers the
    nearest one. Nearest-obstacle cost is the higher of the two."""
    lp = oracle.layer_params()
    occ = np.zeros((48, 48), np.int8)
    for dx, dy in ((6, 6), (8, 3), (3, 8)):
        occ[20 + dy, 20 + dx] = 100
    near = port.costmap_from_occupancy(occ, lp, "nearest")
    prop = port.costmap_from_occupancy(occ, lp, "propagate")
    R, tab = port.inflation_cost_table(lp)
    assert near[20, 20] == tab[6, 6] and prop[20, 20] == tab[8, 3] and tab[6, 6] > tab[8, 3]
    assert (near >= prop).all()


@pytest.mark.parametrize("seed,W,H,kw", [(0, 257, 190, {}), (1, 300, 300, {"inflate_unknown": 1}),
                                        (2, 200, 333, {"resolution": 0.1}), (3, 128, 128, {"trinary_costmap": 0, "gray": True})])
def test_nearest_bounds_propagation(port, seed, W, H, kw):
    gray = kw.pop("gray", False)
    lp = oracle.layer_params(**kw)
    occ = random_occupancy(W, H, seed, gray=gray)
    near = port.costmap_from_occupancy(occ, lp, "nearest")
    prop = port.costmap_from_occupancy(occ, lp, "propagate")
    assert (near >= prop).all()
    assert (near != prop).mean() < 2e-3
    static = port.static_layer(occ, lp)
    assert (near >= static)[static != 255].all()
    # obstacles and cells out of reach are left alone
    assert np.array_equal(near[static == 254], static[static == 254])


def test_rolling_window_of_the_oracle(port):
    """The follower's window (path_follower/config/params.yaml:12-24): away from its rim a window inside the map is
    the crop of the whole-map costmap, beyond the map it is NO_INFORMATION (or free without track_unknown_space),
    and obstacles outside the window do not inflate into it."""
    lp = oracle.layer_params()
    occ = random_occupancy(300, 260, 71, p_lethal=0.005)
    whole = port.costmap_from_occupancy(occ, lp, "nearest")
    R = 10
    win = port.costmap_window_from_occupancy(occ, 40, 30, 200, 200, lp)
    assert np.array_equal(win[R:-R, R:-R], whole[30 + R:230 - R, 40 + R:240 - R])
    out = port.costmap_window_from_occupancy(occ, -50, 100, 120, 80, lp)
    assert (out[:, :50 - R] == 255).all()
    free = port.costmap_window_from_occupancy(occ, -50, 100, 120, 80, oracle.layer_params(track_unknown_space=0))
    assert (free[:, :50 - R] == 0).all()
    one = np.zeros((100, 100), np.int8)
    one[50, 49] = 100  # just left of a window that starts at x = 50
    assert (port.costmap_window_from_occupancy(one, 50, 0, 50, 100, lp) == 0).all()
    assert port.costmap_window_from_occupancy(one, 49, 0, 50, 100, lp)[50, 0] == 254
    # Costmap2D::updateOrigin truncates toward zero on the window's own grid
    assert port.window_origin_cell(5.0, 10.0, 0.05, 0.0) == 0
    assert port.window_origin_cell(12.37, 10.0, 0.05, 0.0) == 147
    assert port.window_origin_cell(0.3, 10.0, 0.05, 0.0) == -94 + 0  # int(-4.7 / 0.05) = -94
    assert port.window_origin_cell(12.37, 10.0, 0.05, -10.0) == 347
