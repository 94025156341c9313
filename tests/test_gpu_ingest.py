"""GPU parity, N2 ingestion: image -> occupancy -> static layer + inflation -> costmap (-> volume refresh),
through the C ABI, bit-exact against oracle/ingest.c."""
import os

import numpy as np
import pytest

from conftest import ROOT, RES
from test_oracle_ingest import random_occupancy

import oracle
from path_planner_b200.capi import LayerParams, PPError

pytestmark = pytest.mark.gpu


def both(**kw):
    """The same layer configuration for the product and for the oracle."""
    return LayerParams.make(**kw), oracle.layer_params(**kw)


@pytest.fixture(scope="module")
def ingest_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "ingest_v1.npz"))


def test_reference_map_asset(handle, ingest_golden):
    g = ingest_golden
    handle.occupancy_from_image(g["image"], float(g["occ_th"]), float(g["free_th"]), bool(g["negate"]))
    assert np.array_equal(handle.occupancy_download(), g["occ"])
    lp, _ = both(resolution=float(g["resolution"]))
    handle.costmap_from_occupancy(lp)
    got = handle.costmap_download()
    assert np.array_equal(got, g["cost_nearest"])
    assert (got >= g["cost_propagate"]).all() and (got != g["cost_propagate"]).sum() <= 7
    handle.occupancy_from_image(g["ramp"], float(g["occ_th"]), float(g["free_th"]), False)
    assert np.array_equal(handle.occupancy_download(), g["ramp_occ"])


@pytest.mark.parametrize("seed,W,H,kw,gray", [
    (0, 257, 190, {}, False),
    (1, 300, 300, {"inflate_unknown": 1}, False),
    (2, 200, 333, {"resolution": 0.1}, False),
    (3, 128, 128, {"trinary_costmap": 0}, True),
    (4, 1000, 37, {"track_unknown_space": 0}, False),
    (5, 129, 65, {"inflation_radius": 0.0}, False),                                  # R = 0: static layer only
    (6, 310, 290, {"resolution": 0.025, "inflation_radius": 0.8}, False),          # R = 32, the limit
    (7, 640, 480, {"cost_scaling_factor": 3.0, "inscribed_radius": 0.3, "inflation_radius": 1.0}, False),
    (8, 96, 96, {"lethal_threshold": 65, "trinary_costmap": 0}, True),
    (9, 31, 17, {}, False),                                                          # smaller than one tile
])
def test_layers_match_oracle(handle, port, seed, W, H, kw, gray):
    lp, olp = both(**kw)
    occ = random_occupancy(W, H, seed, gray=gray)
    handle.occupancy_upload(occ)
    assert np.array_equal(handle.occupancy_download(), occ)
    handle.costmap_from_occupancy(lp)
    want = port.costmap_from_occupancy(occ, olp, "nearest")
    assert np.array_equal(handle.costmap_download(), want)


def test_empty_and_full_maps(handle, port):
    lp, olp = both()
    for fill in (0, 100, -1):
        occ = np.full((70, 150), fill, np.int8)
        handle.occupancy_upload(occ)
        handle.costmap_from_occupancy(lp)
        assert np.array_equal(handle.costmap_download(), port.costmap_from_occupancy(occ, olp, "nearest"))


def test_large_map_against_propagation(handle, port):
    """2048^2: the oracle's brute force is too slow here; costmap_2d's propagation bounds the result from
    below and may differ only in the rare shadowed cells."""
    lp, olp = both()
    occ = random_occupancy(2048, 2048, 21, p_lethal=0.002)
    handle.occupancy_upload(occ)
    handle.costmap_from_occupancy(lp)
    got = handle.costmap_download()
    prop = port.costmap_from_occupancy(occ, olp, "propagate")
    assert (got >= prop).all()
    assert (got != prop).mean() < 1e-3
    # rows 512..640 checked exactly (the brute force on a band with its halo)
    band = port.costmap_from_occupancy(occ[480:672], olp, "nearest")
    assert np.array_equal(got[512:640], band[32:160])


def test_dirty_rect_equals_full_recompute(handle, port):
    lp, olp = both()
    rng = np.random.default_rng(5)
    occ = random_occupancy(700, 500, 31)
    handle.occupancy_upload(occ)
    handle.costmap_from_occupancy(lp)
    for k in range(6):
        w, hgt = int(rng.integers(1, 90)), int(rng.integers(1, 90))
        x0, y0 = int(rng.integers(0, 700 - w)), int(rng.integers(0, 500 - hgt))
        if k == 4:
            x0, y0 = 0, 0              # a corner
        if k == 5:
            x0, y0 = 700 - w, 500 - hgt
        patch = random_occupancy(w, hgt, 100 + k, p_lethal=0.05)
        occ[y0:y0 + hgt, x0:x0 + w] = patch
        handle.occupancy_update_rect(x0, y0, patch)
        if k % 2 == 0:  # two updates merged into one recompute on odd steps
            continue
        handle.costmap_from_occupancy(lp)
        assert np.array_equal(handle.costmap_download(), port.costmap_from_occupancy(occ, olp, "nearest"))
    handle.costmap_from_occupancy(lp)  # nothing pending: a no-op
    assert np.array_equal(handle.costmap_download(), port.costmap_from_occupancy(occ, olp, "nearest"))


def test_map_change_reaches_the_volume(handle, port):
    """occupancy update -> costmap -> pp_se2_volume_refresh gives the volume a full rebuild gives."""
    from path_planner_b200 import capi
    lp, olp = both()
    occ = random_occupancy(384, 256, 41, p_lethal=0.003)
    handle.occupancy_upload(occ)
    handle.costmap_from_occupancy(lp)
    handle.footprint_from_boundary(RES, [0.64, 0.46], True)
    handle.se2_volume_build(72)
    patch = np.zeros((40, 60), np.int8)
    patch[10:30, 20:24] = 100
    occ[100:140, 200:260] = patch
    handle.occupancy_update_rect(200, 100, patch)
    handle.costmap_from_occupancy(lp)
    handle.se2_volume_refresh()
    got = handle.se2_volume_download()
    cost = port.costmap_from_occupancy(occ, olp, "nearest")
    assert np.array_equal(handle.costmap_download(), cost)
    fresh = capi.Handle(0)
    try:
        fresh.costmap_upload(cost)
        fresh.footprint_from_boundary(RES, [0.64, 0.46], True)
        fresh.se2_volume_build(72)
        assert np.array_equal(got, fresh.se2_volume_download())
    finally:
        fresh.close()


def test_parameter_change_recomputes_everything(handle, port):
    occ = random_occupancy(200, 160, 51)
    handle.occupancy_upload(occ)
    for kw in ({}, {"cost_scaling_factor": 4.0}, {"inflation_radius": 0.3}, {"track_unknown_space": 0}):
        lp, olp = both(**kw)
        handle.costmap_from_occupancy(lp)
        assert np.array_equal(handle.costmap_download(), port.costmap_from_occupancy(occ, olp, "nearest"))


def test_errors(handle):
    lp, _ = both()
    with pytest.raises(PPError) as e:
        handle.costmap_from_occupancy(lp)  # nothing uploaded
    assert e.value.status == 2
    occ = np.zeros((64, 64), np.int8)
    handle.occupancy_upload(occ)
    with pytest.raises(PPError) as e:
        handle.costmap_from_occupancy(LayerParams.make(resolution=0.01, inflation_radius=0.5))  # 50 cells
    assert e.value.status == 4
    with pytest.raises(PPError):
        handle.costmap_from_occupancy(LayerParams.make(resolution=0.0))
    with pytest.raises(PPError):
        handle.occupancy_update_rect(60, 60, np.zeros((8, 8), np.int8))  # sticks out


# ---- the follower's rolling window (path_follower/config/params.yaml:12-24) ---------------------------

@pytest.mark.parametrize("origin,win,kw", [
    ((37, 21), (200, 200), {}),                                 # inside the map: the follower's 10 m x 10 m at 0.05
    ((-60, 150), (200, 200), {}),                               # sticks out to the left: unknown beyond the map
    ((310, 260), (200, 200), {"track_unknown_space": 0}),       # over the far corner, unknown taken as free
    ((-500, -500), (64, 48), {}),                               # entirely off the map
    ((100, 90), (131, 77), {"inflate_unknown": 1}),             # ragged size
    ((-10, -10), (420, 380), {"inflation_radius": 0.8}),        # larger than the map
])
def test_rolling_window_matches_oracle(handle, port, origin, win, kw):
    from path_planner_b200 import capi
    lp, olp = both(**kw)
    occ = random_occupancy(400, 350, 61)
    src = capi.Handle(0)  # the global map's handle
    try:
        src.occupancy_upload(occ)
        handle.costmap_window_from_occupancy(src, origin[0], origin[1], win[0], win[1], lp)
        got = handle.costmap_download()
        assert got.shape == (win[1], win[0])
        assert np.array_equal(got, port.costmap_window_from_occupancy(occ, origin[0], origin[1], win[0], win[1], olp))
    finally:
        src.close()


def test_rolling_window_follows_the_robot_and_feeds_the_obstacle_search(handle, port):
    """A window moved along a path on ONE handle that also holds the map; every window equals the oracle's, a window
    far from the edges equals the crop of the whole-map costmap away from its rim, and the obstacle extraction of
    the follower (K3a) runs on the window."""
    lp, olp = both()
    occ = random_occupancy(600, 500, 62, p_lethal=0.004)
    handle.occupancy_upload(occ)
    handle.costmap_from_occupancy(lp)
    whole = handle.costmap_download()
    R = 10  # ceil(0.5 / 0.05)
    for robot in [(5.0, 5.0), (12.37, 9.81), (24.99, 20.02), (0.3, 24.0)]:
        ox = handle.L.pp_window_origin_cell(robot[0], 10.0, 0.05, 0.0)
        oy = handle.L.pp_window_origin_cell(robot[1], 10.0, 0.05, 0.0)
        assert (ox, oy) == (port.window_origin_cell(robot[0], 10.0, 0.05, 0.0), port.window_origin_cell(robot[1], 10.0, 0.05, 0.0))
        handle.costmap_window_from_occupancy(handle, ox, oy, 200, 200, lp)
        got = handle.costmap_download()
        assert np.array_equal(got, port.costmap_window_from_occupancy(occ, ox, oy, 200, 200, olp))
        if ox >= 0 and oy >= 0 and ox + 200 <= 600 and oy + 200 <= 500:
            assert np.array_equal(got[R:-R, R:-R], whole[oy + R:oy + 200 - R, ox + R:ox + 200 - R])
        # the window's world origin is (ox, oy) cells from the map's; the robot sits in its middle
        origin_w = (ox * 0.05, oy * 0.05)
        pose = (robot[0], robot[1], 0.4)
        pts = handle.obstacles_extract(0.05, origin_w, pose, 4.5, 1.5)
        want = port.extract_obstacles(got, 0.05, origin_w, pose, 4.5, 1.5)
        assert np.array_equal(pts, want)
    # back to the whole map on the same handle
    handle.costmap_from_occupancy(lp)
    assert np.array_equal(handle.costmap_download(), whole)


def test_rolling_window_errors(handle):
    from path_planner_b200 import capi
    lp, _ = both()
    with pytest.raises(PPError) as e:
        handle.costmap_window_from_occupancy(handle, 0, 0, 200, 200, lp)  # no occupancy grid
    assert e.value.status == 2
    handle.occupancy_upload(np.zeros((64, 64), np.int8))
    with pytest.raises(PPError):
        handle.costmap_window_from_occupancy(handle, 0, 0, 0, 200, lp)
    with pytest.raises(PPError):
        handle.costmap_window_from_occupancy(handle, 0, 0, 40000, 10, lp)
