"""GPU parity: K3 TEB candidate extraction and radius / top-K association."""
import numpy as np
import pytest

from path_planner_b200.synth import make_costmap, make_teb_scene

pytestmark = pytest.mark.gpu


def assert_topk_equal(got, want):
    idx, d2, cnt = got
    widx, wd2, wcnt = want[0], want[1], want[2]
    assert np.array_equal(cnt, wcnt)
    assert np.array_equal(d2, wd2)  # doubles, bit-exact
    diff = idx != widx
    for q, j in zip(*np.nonzero(diff)):  # only permutations among equal distances are tolerated
        assert j < cnt[q] and (wd2[q, :cnt[q]] == d2[q, j]).sum() > 1, (q, j)


def test_radius_topk_matches_golden(handle, golden):
    got = handle.radius_topk(golden["teb_pts"], golden["teb_queries"], 0.9574, 10)
    assert_topk_equal(got, (golden["teb_idx"], golden["teb_d2"], golden["teb_cnt"]))


@pytest.mark.parametrize("N,Q,K,r", [(10000, 100, 10, 0.9574), (10000, 100, 5, 0.9574), (500, 7, 64, 0.5),
                                     (10000, 3, 10, 1e9), (1, 1, 1, 1e9), (5000, 64, 10, 1e-9)])
def test_radius_topk_vs_oracle(handle, port, N, Q, K, r):
    pts, poses = make_teb_scene(N, Q, seed=N + Q)
    got = handle.radius_topk(pts, poses, r, K)
    assert_topk_equal(got, port.radius_topk(pts, poses, r, K))


def test_radius_topk_ties_and_strictness(handle, port):
    # points on a lattice: many exactly equal distances, and points exactly at d2 == radius_arg
    g = np.arange(-5, 6, dtype=np.float64)
    pts = np.stack(np.meshgrid(g, g), -1).reshape(-1, 2)
    q = np.array([[0.0, 0.0], [0.5, 0.5], [5.0, 5.0]])
    for r in (1.0, 2.0, 4.0, 0.5):  # d2 == r must be excluded (strict <)
        got = handle.radius_topk(pts, q, r, 8)
        want = port.radius_topk(pts, q, r, 8)
        assert_topk_equal(got, want)
        assert (got[1][got[1] > 0] < r).all()


def test_three_rounds_with_moving_poses(handle, port):
    """One replan = 1 extraction + 3 query rounds with the poses moved in between (teb_optimizer.cpp:90-108)."""
    pts, poses = make_teb_scene(10000, 100, seed=1)
    rng = np.random.default_rng(2)
    handle.radius_topk(pts, poses, 0.9574, 10)
    for _ in range(3):
        poses = poses + rng.uniform(-0.02, 0.02, poses.shape)
        got = handle.radius_topk(None, poses, 0.9574, 10)  # points stay resident on the device
        assert_topk_equal(got, port.radius_topk(pts, poses, 0.9574, 10))


@pytest.mark.parametrize("shape,pose", [((200, 200), (5.0, 5.0, 0.7)), ((200, 200), (2.0, 8.0, -2.5)), ((77, 130), (1.0, 3.0, 3.1)),
                                        ((1, 1), (0.0, 0.0, 0.0)), ((640, 480), (16.0, 12.0, 1.3))])
def test_extract_matches_oracle(handle, port, shape, pose):
    W, H = shape
    cost = make_costmap(W, H, seed=W + H, coverage=0.1) if W > 8 else np.full((H, W), 254, np.uint8)
    handle.costmap_upload(cost)
    for (mfr, behind) in [(4.5, 1.5), (1.0, 0.3), (100.0, 0.0)]:
        got = handle.obstacles_extract(0.05, (0.0, 0.0), pose, mfr, behind)
        want = port.extract_obstacles(cost, 0.05, (0.0, 0.0), pose, mfr, behind)
        assert got.shape == want.shape
        assert np.array_equal(got, want)


def test_extract_then_query_on_device(handle, port):
    cost = make_costmap(200, 200, seed=10, coverage=0.08)
    pose = (5.0, 5.0, 0.3)
    handle.costmap_upload(cost)
    pts = handle.obstacles_extract(0.05, (0.0, 0.0), pose, 4.5, 1.5)
    s = np.linspace(0, 3, 33)
    poses = np.stack([5.0 + s * np.cos(0.3), 5.0 + s * np.sin(0.3)], 1)
    got = handle.radius_topk(None, poses, 0.9574, 10)
    assert_topk_equal(got, port.radius_topk(pts, poses, 0.9574, 10))


def test_empty_extraction(handle):
    handle.costmap_upload(np.zeros((50, 50), np.uint8))
    pts = handle.obstacles_extract(0.05, (0.0, 0.0), (1.0, 1.0, 0.0), 4.5, 1.5)
    assert pts.shape == (0, 2)
    idx, d2, cnt = handle.radius_topk(None, np.zeros((4, 2)), 0.9574, 10)
    assert (cnt == 0).all()


@pytest.mark.parametrize("n_in", [3500, 9000])
def test_dense_cluster_many_hits_per_query(handle, port, n_in):
    """More hits inside the radius than the kernel buffers between two reductions (several reduction passes per
    query, hits arriving from every warp): the keep-first-K result must not depend on the buffering."""
    rng = np.random.default_rng(n_in)
    ang, rad = rng.uniform(0, 2 * np.pi, n_in), 0.9 * np.sqrt(rng.uniform(0, 1, n_in))
    cluster = np.stack([1.0 + rad * np.cos(ang), -2.0 + rad * np.sin(ang)], 1)
    far = rng.uniform(-50, 50, (2000, 2))
    pts = np.ascontiguousarray(rng.permutation(np.concatenate([cluster, far])))
    q = np.array([[1.0, -2.0], [1.3, -2.2], [30.0, 30.0], [0.2, -1.9]])
    for K in (10, 64):
        got = handle.radius_topk(pts, q, 0.9574, K)
        want = port.radius_topk(pts, q, 0.9574, K)
        assert_topk_equal(got, want)
    assert port.radius_topk(pts, q, 0.9574, 1)[3][0] > 3072  # `found` of the first query: above the kernel's buffer


@pytest.mark.parametrize("origin", [(0.0, 0.0), (-7.3, 2.1)])
def test_column_pruned_query_equals_unpruned(handle, port, origin):
    """Points left on the device by the extraction are column-ordered; queries scan only the columns in reach.
    Same result as the full scan, for poses inside, at the edges of and outside the map."""
    cost = make_costmap(300, 220, seed=21, coverage=0.15)
    cost[:, 0] = 254  # obstacles in the first and the last column
    cost[:, -1] = 254
    handle.costmap_upload(cost)
    pose = (origin[0] + 7.0, origin[1] + 5.0, 0.4)
    pts = handle.obstacles_extract(0.05, origin, pose, 100.0, 0.0)
    assert len(pts) > 5000
    rng = np.random.default_rng(3)
    q = np.concatenate([rng.uniform(-1.0, 16.0, (60, 2)) + np.array(origin), np.array([[0.0, 5.0], [15.0, 5.0], [-30.0, 0.0], [40.0, 3.0]]) + np.array(origin)])
    for r in (0.9574, 0.01, 30.0):
        got = handle.radius_topk(None, q, r, 10)
        assert_topk_equal(got, port.radius_topk(pts, q, r, 10))


def test_extract_candidates_exactly_at_the_behind_range(handle, port):
    """teb_follower.cpp:158-161 compares glibc's hypot with obstacle_behind_ignore_range. Obstacles at distance
    5 cells x 1 m = exactly the range, and one ulp either side of it, behind the robot: the kernel hands such
    candidates to the host's hypot instead of trusting the device's."""
    cost = np.zeros((41, 41), np.uint8)
    for (dx, dy) in [(-3, 4), (-4, 3), (-5, 0), (-3, -4), (-4, -3), (0, 5), (3, 4), (-6, 8), (-8, 6), (-10, 0)]:
        cost[20 + dy, 20 + dx] = 254
    handle.costmap_upload(cost)
    pose = (20.5, 20.5, 0.0)  # centre of cell (20, 20) at 1 m resolution, looking along +x
    for R in (5.0, np.nextafter(5.0, 0.0), np.nextafter(5.0, 10.0), 10.0, np.nextafter(10.0, 0.0), 4.999, 0.0):
        got = handle.obstacles_extract(1.0, (0.0, 0.0), pose, 100.0, float(R))
        want = port.extract_obstacles(cost, 1.0, (0.0, 0.0), pose, 100.0, float(R))
        assert got.shape == want.shape and np.array_equal(got, want), R
    # the counts really change across the boundary (the case is not vacuous)
    n_at = len(port.extract_obstacles(cost, 1.0, (0.0, 0.0), pose, 100.0, 5.0))
    n_below = len(port.extract_obstacles(cost, 1.0, (0.0, 0.0), pose, 100.0, float(np.nextafter(5.0, 0.0))))
    assert n_at > n_below
