"""GPU parity: K1 SE2 footprint-collision volume through the C ABI, bit-exact against the golden
vectors (reference's own code) and the oracle on seeded inputs."""
import numpy as np
import pytest

from conftest import FOOTPRINTS, OFFSETS, RES
from path_planner_b200.synth import make_costmap

pytestmark = pytest.mark.gpu


def build(handle, cost, fname, n_theta=72, off=(0.5, 0.5, 0.0), t0=0, t1=None, mode=0):
    b, hole = FOOTPRINTS[fname]
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, b, hole)
    handle.se2_volume_set_mode(mode)
    handle.se2_volume_build(n_theta, *off, t_begin=t0, t_end=t1)
    return handle.se2_volume_download(t0, t1)


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("fname,oname", [(f, o) for f in FOOTPRINTS for o in OFFSETS if f == "rect_hole" or o == "centre"])
def test_volume_matches_golden(handle, golden, fname, oname, mode):
    v = build(handle, golden["cost"], fname, 72, OFFSETS[oname], mode=mode)
    assert np.array_equal(v, golden[f"vol_{fname}_{oname}_72"])


def test_volume_144_bins_matches_golden(handle, golden):
    v = build(handle, golden["cost"], "rect_hole", 144)
    bins = golden["vol_rect_hole_centre_144_bins"]
    assert np.array_equal(v[bins], golden["vol_rect_hole_centre_144"])


def test_footprint_points_match_reference(handle, golden):
    for fname, (b, hole) in FOOTPRINTS.items():
        handle.footprint_from_boundary(RES, b, hole)
        assert np.array_equal(handle.footprint_get(), golden[f"fp_{fname}"])


@pytest.mark.parametrize("shape", [(1, 1), (3, 2), (31, 33), (131, 77), (200, 150), (257, 64)])
@pytest.mark.parametrize("mode", [0, 1])
def test_volume_ragged_sizes_vs_oracle(handle, port, shape, mode):
    W, H = shape
    cost = make_costmap(W, H, seed=W * 1000 + H, coverage=0.1) if W > 8 else np.full((H, W), 254, np.uint8)
    for fname in ("rect_hole", "box_hole"):
        for off in OFFSETS.values():
            v = build(handle, cost, fname, 72, off, mode=mode)
            b, hole = FOOTPRINTS[fname]
            o = port.se2_volume(cost, port.footprint_points(RES, b, hole), 72, *off)
            assert np.array_equal(v, o), (shape, fname, off)


@pytest.mark.parametrize("fill", [0, 1, 252, 253, 254, 255])
def test_volume_constant_maps(handle, port, fill):
    cost = np.full((70, 90), fill, np.uint8)
    v = build(handle, cost, "rect_hole")
    o = port.se2_volume(cost, port.footprint_points(RES, *FOOTPRINTS["rect_hole"]), 72)
    assert np.array_equal(v, o)
    # every pose has at least one in-map sample on a 90x70 map, so the volume is constant too
    assert (v == fill).all()


def test_volume_partial_bins_and_arbitrary_offsets(handle, port):
    cost = make_costmap(150, 90, seed=5)
    offs = np.array([[0, 0], [3, -2], [-7, 1], [0, 9], [4, 4], [4, 4], [-1, -1]], np.int32)  # with a duplicate
    handle.costmap_upload(cost)
    handle.footprint_set(offs)
    handle.se2_volume_build(72, 0.5, 0.5, 0.0, 18, 27)
    v = handle.se2_volume_download(18, 27)
    o = port.se2_volume(cost, offs, 72, 0.5, 0.5, 0.0, t0=18, t1=27)
    assert np.array_equal(v, o)


def test_dirty_region_refresh_equals_rebuild(handle, port):
    cost = make_costmap(300, 200, seed=6)
    pts = port.footprint_points(RES, *FOOTPRINTS["rect_hole"])
    v0 = build(handle, cost, "rect_hole")
    assert np.array_equal(v0, port.se2_volume(cost, pts, 72))
    rng = np.random.default_rng(0)
    new = cost.copy()
    for (x0, y0, w, h) in [(40, 50, 30, 20), (250, 170, 50, 30), (0, 0, 10, 10)]:
        patch = rng.choice(np.array([0, 0, 0, 100, 253, 254, 255], np.uint8), size=(h, w))
        new[y0:y0 + h, x0:x0 + w] = patch
        handle.costmap_update_rect(x0, y0, patch)
    handle.se2_volume_refresh()
    v1 = handle.se2_volume_download()
    assert np.array_equal(handle.costmap_download(), new)
    assert np.array_equal(v1, port.se2_volume(new, pts, 72))
    handle.se2_volume_refresh()  # nothing dirty: a no-op
    assert np.array_equal(handle.se2_volume_download(), v1)


def test_fast_and_generic_paths_agree_on_large_map(handle):
    cost = make_costmap(1100, 700, seed=9)
    a = build(handle, cost, "rect_hole", mode=0)
    st = handle.se2_volume_stats()
    b = build(handle, cost, "rect_hole", mode=1)
    assert np.array_equal(a, b), st


def test_full_size_volume_sampled_against_oracle(handle, port):
    """BASELINE config 3 (4096x4096x72): too big for the CPU oracle as a whole, so check (a) a seeded
    sample of cells against the oracle's per-pose evaluation, (b) two full row bands, (c) that
    rebuilding is deterministic, (d) that collision flags never appear on an all-free map."""
    W = H = 4096
    cost = make_costmap(W, H, seed=1)
    b, hole = FOOTPRINTS["rect_hole"]
    pts = port.footprint_points(RES, b, hole)
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, b, hole)
    handle.se2_volume_build(72)
    rng = np.random.default_rng(3)
    bin_sz = np.float32(2.0) * np.float32(np.pi) / np.float32(72)
    for t in (0, 1, 17, 18, 35, 36, 54, 71):
        plane = handle.se2_volume_download(t, t + 1)[0]
        xs = rng.integers(0, W, 3000)
        ys = rng.integers(0, H, 3000)
        # include the map corners and power-of-two columns/rows where float spacing changes
        xs[:8] = [0, 1, W - 1, W - 2, 2047, 2048, 1023, 1024]
        ys[:8] = [0, W - 1, 0, H - 1, 2048, 2047, 1024, 1023]
        theta = np.float32(np.float32(t) * bin_sz)
        poses = np.stack([xs.astype(np.float32) + np.float32(0.5), ys.astype(np.float32) + np.float32(0.5),
                          np.full(3000, theta, np.float32)], 1)
        c, _, _ = port.footprint_cost_batch(cost, pts, poses)
        assert np.array_equal(plane[ys, xs], c.astype(np.uint8)), t
        if t in (17, 54):
            for y0 in (0, 2040):
                band = port.se2_volume(cost, pts, 72, t0=t, t1=t + 1, y0=y0, y1=y0 + 16)
                assert np.array_equal(plane[y0:y0 + 16], band[0]), (t, y0)
    first = handle.se2_volume_download(10, 12).copy()
    handle.se2_volume_build(72)
    assert np.array_equal(handle.se2_volume_download(10, 12), first)
    handle.costmap_upload(np.zeros((H, W), np.uint8))
    handle.se2_volume_build(72, t_begin=0, t_end=2)
    assert not handle.se2_volume_download(0, 2).any()


def test_error_paths(handle):
    from path_planner_b200.capi import PPError
    with pytest.raises(PPError):
        handle.se2_volume_build(72)  # no costmap
    handle.costmap_upload(np.zeros((8, 8), np.uint8))
    with pytest.raises(PPError):
        handle.se2_volume_build(72)  # no footprint
    handle.footprint_from_boundary(RES, [0.64, 0.46], True)
    with pytest.raises(PPError):
        handle.se2_volume_build(72, t_begin=5, t_end=80)
    with pytest.raises(PPError):
        handle.footprint_from_boundary(RES, [1.0, 2.0, 3.0], True)  # 3 numbers: the reference asserts
    with pytest.raises(PPError):
        handle.se2_volume_refresh()  # nothing built yet


def test_arbitrary_footprints_through_both_paths(handle, port):
    """pp_footprint_set with random offset lists (duplicates, lopsided shapes, single points): the
    planner has to find exact covers for sets no library was designed for."""
    cost = make_costmap(320, 192, seed=31)
    rng = np.random.default_rng(7)
    handle.costmap_upload(cost)
    for trial in range(6):
        n = int(rng.integers(1, 70))
        offs = rng.integers(-9, 10, size=(n, 2)).astype(np.int32)
        if trial == 0:
            offs = np.array([[0, 0]], np.int32)
        if trial == 1:
            offs = np.concatenate([offs, offs[:5]])  # duplicates
        want = port.se2_volume(cost, offs, 72, 0.5, 0.5, 0.0)
        for mode in (0, 1):
            handle.footprint_set(offs)
            handle.se2_volume_set_mode(mode)
            handle.se2_volume_build(72)
            assert np.array_equal(handle.se2_volume_download(), want), (trial, mode)
    handle.se2_volume_set_mode(0)


@pytest.mark.parametrize("n_theta,off", [(7, (0.5, 0.5, 0.0)), (144, (0.25, 0.75, 0.5)), (1, (0.0, 0.0, 0.0)), (72, (-0.3, 0.9, 0.0))])
def test_odd_bin_counts_and_offsets(handle, port, n_theta, off):
    cost = make_costmap(260, 130, seed=32)
    pts = port.footprint_points(RES, [0.64, 0.46], True)
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, [0.64, 0.46], True)
    handle.se2_volume_build(n_theta, *off)
    assert np.array_equal(handle.se2_volume_download(), port.se2_volume(cost, pts, n_theta, *off))


def test_large_footprint_and_coarse_resolution(handle, port):
    """reach 22 cells (beyond the fast path's halo limit: everything goes to the generic kernel)
    and the 0.1 m resolution of the shipped map (28-point rim)."""
    cost = make_costmap(200, 160, seed=33)
    handle.costmap_upload(cost)
    for res, b, hole in [(0.05, [1.6, 1.2], True), (0.1, [0.64, 0.46], True), (0.05, [1.0], False)]:
        pts = port.footprint_points(res, b, hole)
        handle.footprint_from_boundary(res, b, hole)
        handle.se2_volume_build(72, t_begin=30, t_end=40)
        got = handle.se2_volume_download(30, 40)
        assert np.array_equal(got, port.se2_volume(cost, pts, 72, t0=30, t1=40)), (res, b)


def test_download_rows_is_a_slice_of_the_volume(handle):
    cost = make_costmap(200, 150, seed=8)
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, [0.64, 0.46], True)
    handle.se2_volume_build(72)
    vol = handle.se2_volume_download()
    for (y0, y1, t0, t1) in ((0, 150, 0, 72), (32, 96, 0, 72), (149, 150, 10, 11), (16, 48, 36, 72)):
        assert np.array_equal(handle.se2_volume_download_rows(y0, y1, t0, t1), vol[t0:t1, y0:y1])
    from path_planner_b200.capi import PPError
    with pytest.raises(PPError):
        handle.se2_volume_download_rows(100, 151)


def test_twin_bins_equal_their_own_evaluation(handle, port):
    """Bins t and t + n/2 of a centrally symmetric footprint share one evaluation in the fast kernel; a build
    of the upper half alone (no twin in the launch) must give the same bytes, and both must match the oracle."""
    cost = make_costmap(448, 200, seed=12)
    pts = port.footprint_points(RES, [0.64, 0.46], True)
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, [0.64, 0.46], True)
    handle.se2_volume_build(72)
    whole = handle.se2_volume_download()
    handle.se2_volume_build(72, t_begin=36, t_end=72)
    upper = handle.se2_volume_download(36, 72)
    assert np.array_equal(whole[36:], upper)
    want = port.se2_volume(cost, pts, 72, t0=30, t1=42)
    assert np.array_equal(whole[30:42], want)


@pytest.mark.parametrize("boundary", [[0.5, 0.4], [0.6, 0.5], [0.64, 0.46], [0.8, 0.6], [0.9, 0.6], [1.0, 0.6], [1.1, 0.6],
                                      [1.2, 0.7], [1.3, 0.7]])
def test_every_halo_instantiation(handle, port, boundary):
    """Footprint reaches 7 ... 15 cells: one per compiled (pitch, halo) variant of the tiled kernel."""
    cost = make_costmap(320, 208, seed=int(boundary[0] * 100))
    pts = port.footprint_points(RES, boundary, True)
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, boundary, True)
    handle.se2_volume_build(72)
    stats = handle.se2_volume_stats()
    assert stats["fast_tiles"] > 0  # the tiled kernel did take part
    assert np.array_equal(handle.se2_volume_download(), port.se2_volume(cost, pts, 72))


def test_build_of_a_scattered_bin_set(handle, golden):
    """pp_se2_volume_build_bins: any set of heading bins (here pairs t, t + 36 as a heading-sharded rank owns
    them, a lone bin, and bins given out of order and twice); the other planes of the volume stay untouched."""
    cost = golden["cost"]
    want = golden["vol_rect_hole_centre_72"]
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(0.05, *FOOTPRINTS["rect_hole"])
    handle.se2_volume_build(72)
    handle.costmap_upload(np.zeros_like(cost))
    handle.se2_volume_build(72)  # all planes now hold the empty map's volume
    handle.costmap_upload(cost)
    bins = [40, 4, 5, 41, 17, 4, 70, 34]
    handle.se2_volume_build_bins(72, bins)
    got = handle.se2_volume_download()
    for t in range(72):
        if t in bins:
            assert np.array_equal(got[t], want[t]), t
        else:
            assert not got[t].any(), t
    # a refresh after a costmap change recomputes exactly these bins
    patch = np.full((9, 11), 254, np.uint8)
    handle.costmap_update_rect(20, 30, patch)
    handle.se2_volume_refresh()
    c2 = cost.copy()
    c2[30:39, 20:31] = 254
    handle.costmap_upload(c2)
    got2 = handle.se2_volume_download()
    handle.se2_volume_build(72)
    full = handle.se2_volume_download()
    for t in set(bins):
        assert np.array_equal(got2[t], full[t]), t


def test_short_reach_is_refused_not_clamped(monkeypatch):
    """Both kernels stage reach + 2 cells around a tile. If the reach were ever short of a sample offset the
    result would be wrong, so such a configuration is refused (PP_ERR_RANGE) by either kernel variant. The
    bound cannot be broken through the ABI (the reach is derived from the footprint); a debug knob shortens it."""
    from path_planner_b200 import capi
    cost = make_costmap(200, 150, seed=2)
    for mode in (0, 1):
        monkeypatch.setenv("PP_DEBUG_REACH_BIAS", "-4")
        h = capi.Handle(0)
        h.costmap_upload(cost)
        h.footprint_from_boundary(RES, *FOOTPRINTS["rect_hole"])
        monkeypatch.delenv("PP_DEBUG_REACH_BIAS")
        h.se2_volume_set_mode(mode)
        with pytest.raises(capi.PPError) as ei:
            h.se2_volume_build(72)
        assert ei.value.status == 4  # PP_ERR_RANGE
        h.close()


def test_new_entry_points_reject_bad_arguments(handle):
    from path_planner_b200 import capi
    cost = make_costmap(160, 96, seed=4)
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, *FOOTPRINTS["rect_hole"])
    with pytest.raises(capi.PPError):
        handle.se2_volume_build_bins(72, [3, 72])          # bin outside [0, n_theta)
    with pytest.raises(capi.PPError) as ei:
        handle.se2_volume_build_sharded(72, 3)              # no communicator
    assert ei.value.status == 2  # PP_ERR_STATE
    handle.se2_volume_build_bins(72, [])                    # an empty set is a no-op
    handle.se2_volume_build_bins(72, [5])
    assert handle.se2_volume_download(5, 6).any()


@pytest.mark.parametrize("shape", [(96, 80), (200, 150), (256, 64), (400, 400)])
@pytest.mark.parametrize("fname", ["rect_hole", "circle_hole", "rect_solid"])
def test_32_row_tiles_match_the_port(handle, port, monkeypatch, shape, fname):
    """Large builds use 128 x 32 tiles (the planner picks them from ~1000 tiles on); forced here on small, ragged maps
    so that partial tiles, mixed tiles at the first columns / rows and the dirty-region refresh run with them too."""
    monkeypatch.setenv("PP_SE2_TILE_ROWS", "32")
    W, H = shape
    cost = make_costmap(W, H, seed=W + H, coverage=0.1)
    b, hole = FOOTPRINTS[fname]
    pts = port.footprint_points(RES, b, hole)
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, b, hole)
    handle.se2_volume_build(72)
    assert np.array_equal(handle.se2_volume_download(), port.se2_volume(cost, pts, 72))
    c2 = cost.copy()
    c2[H // 3:H // 3 + 9, W // 2:W // 2 + 13] = 254
    handle.costmap_update_rect(W // 2, H // 3, np.ascontiguousarray(c2[H // 3:H // 3 + 9, W // 2:W // 2 + 13]))
    handle.se2_volume_refresh()
    assert np.array_equal(handle.se2_volume_download(), port.se2_volume(c2, pts, 72))
