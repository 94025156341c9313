"""GPU parity: K0 upload / dirty rect and K1b batched footprint cost at arbitrary poses."""
import numpy as np
import pytest

from conftest import FOOTPRINTS, RES
from path_planner_b200.synth import make_costmap

pytestmark = pytest.mark.gpu


def test_upload_roundtrip_and_rect(handle):
    cost = make_costmap(333, 211, seed=2)
    handle.costmap_upload(cost)
    assert np.array_equal(handle.costmap_download(), cost)
    patch = np.arange(7 * 13, dtype=np.uint8).reshape(7, 13)
    handle.costmap_update_rect(320, 204, patch)
    cost[204:211, 320:333] = patch
    assert np.array_equal(handle.costmap_download(), cost)
    from path_planner_b200.capi import PPError
    with pytest.raises(PPError):
        handle.costmap_update_rect(330, 0, patch)  # sticks out


def test_upload_from_pinned_memory(handle):
    from path_planner_b200.capi import PinnedBuffer
    cost = make_costmap(128, 64, seed=3)
    buf = PinnedBuffer(cost.size)
    view = buf.array.reshape(cost.shape)
    view[:] = cost
    handle.costmap_upload(view, pinned=True)
    assert np.array_equal(handle.costmap_download(), cost)
    buf.free()


def test_batch_matches_golden(handle, golden):
    handle.costmap_upload(golden["cost"])
    handle.footprint_from_boundary(RES, [0.64, 0.46], True)
    c, f = handle.footprint_cost_batch(golden["batch_poses"])
    assert np.array_equal(c, golden["batch_cost"])
    assert np.array_equal(f, golden["batch_flag"])
    assert np.array_equal(f, golden["batch_collision"])


@pytest.mark.parametrize("fname", list(FOOTPRINTS))
def test_batch_random_poses_vs_oracle(handle, port, fname):
    cost = make_costmap(500, 300, seed=12)
    b, hole = FOOTPRINTS[fname]
    rng = np.random.default_rng(8)
    n = 20000
    poses = np.stack([rng.uniform(-15, 515, n), rng.uniform(-15, 315, n), rng.uniform(-10, 10, n)], 1).astype(np.float32)
    poses[:4] = [[0, 0, 0], [-0.0, -0.0, 0], [499.99, 299.99, 3.14159], [-1e-7, 5, 1.0]]
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, b, hole)
    c, f = handle.footprint_cost_batch(poses)
    oc, of, ok = port.footprint_cost_batch(cost, port.footprint_points(RES, b, hole), poses)
    assert np.array_equal(c, oc)
    assert np.array_equal(f, of)


def test_batch_agrees_with_volume_at_bin_poses(handle):
    """The volume is by definition the batched function at the canonical pose of each cell."""
    cost = make_costmap(160, 120, seed=4)
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, [0.64, 0.46], True)
    handle.se2_volume_build(72)
    v = handle.se2_volume_download()
    rng = np.random.default_rng(1)
    t = rng.integers(0, 72, 5000)
    x = rng.integers(0, 160, 5000)
    y = rng.integers(0, 120, 5000)
    bin_sz = np.float32(2.0) * np.float32(np.pi) / np.float32(72)
    poses = np.stack([x.astype(np.float32) + np.float32(0.5), y.astype(np.float32) + np.float32(0.5),
                      t.astype(np.float32) * bin_sz], 1).astype(np.float32)
    c, f = handle.footprint_cost_batch(poses)
    assert np.array_equal(c.astype(np.uint8), v[t, y, x])
    assert np.array_equal(f.astype(bool), v[t, y, x] >= 253)


def test_batch_empty(handle):
    handle.costmap_upload(np.zeros((4, 4), np.uint8))
    handle.footprint_from_boundary(RES, [0.35], False)
    c, f = handle.footprint_cost_batch(np.zeros((0, 3), np.float32))
    assert len(c) == 0 and len(f) == 0
