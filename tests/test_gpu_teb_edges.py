"""GPU parity, N3: batched EdgeObstacle errors and numeric Jacobians through the C ABI against the
fixture generated from the reference's own code and against the oracle on fresh inputs.

Tolerances (floating point, double): errors 1e-12 absolute -- the device hypot/sincos may differ from
glibc's in the last bit; Jacobians 1e-6 absolute -- g2o's central difference divides error differences
by 2e-9, so a last-bit difference (1e-16) in an error becomes 1e-7 in a Jacobian entry, in the reference
itself as much as here."""
import os

import numpy as np
import pytest

from conftest import ROOT

from path_planner_b200.capi import ObstacleEdgeParams, PPError

pytestmark = pytest.mark.gpu

ERR_TOL = 1e-12
JAC_TOL = 1e-6


@pytest.fixture(scope="module")
def edges_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "teb_edges_v1.npz"))


@pytest.mark.parametrize("name", ("rect", "circle", "pentagon"))
def test_matches_reference_fixture(handle, edges_golden, name):
    g = edges_golden
    K = int(g["K"])
    idx, d2, cnt = handle.radius_topk(g["pts"], g["poses"][:, :2], float(g["radius_arg"]), K)
    assert np.array_equal(cnt, g["cnt"])
    used = np.arange(K)[None, :] < cnt[:, None]
    assert np.array_equal(idx[used], g["idx"][used])
    ri, rc = handle.robot_model_set(g[f"robot_{name}"])
    assert (ri, rc) == tuple(g[f"radii_{name}"])
    err, jac = handle.obstacle_edges(g["poses"], K, ObstacleEdgeParams.make(*g["info"]))
    assert np.abs(err - g[f"err_{name}"]).max() <= ERR_TOL
    assert np.abs(jac - g[f"jac_{name}"]).max() <= JAC_TOL
    assert np.array_equal(err == 0, g[f"err_{name}"] == 0)  # same penalty branch everywhere
    err_only, none = handle.obstacle_edges(g["poses"], K, ObstacleEdgeParams.make(*g["info"]), jacobian=False)
    assert none is None and np.array_equal(err_only, err)


def test_fresh_inputs_and_partial_association(handle, port):
    rng = np.random.default_rng(17)
    pts = rng.uniform(-3, 3, (3000, 2))
    poses = np.concatenate([rng.uniform(-2.5, 2.5, (100, 2)), rng.uniform(-7, 7, (100, 1))], axis=1)
    for K, r_arg, robot, info in ((10, 0.9574, [0.72, 0.50], [0.1, 0.05, 0.2, 0.1]),     # shipped defaults
                                  (5, 0.05, [0.3], [0.1, 0.05, 0.2, 0.1]),                 # few hits: many unused slots
                                  (64, 3.0, [0.5, 0.0, -0.3, 0.3, -0.3, -0.3], [0.3, 0.0, 0.6, 0.2])):
        idx, d2, cnt = handle.radius_topk(pts, poses[:, :2], r_arg, K)
        handle.robot_model_set(robot)
        err, jac = handle.obstacle_edges(poses, K, ObstacleEdgeParams.make(*info))
        e0, j0 = port.obstacle_edges(poses, pts, idx, cnt, robot, info)
        assert np.abs(err - e0).max() <= ERR_TOL
        assert np.abs(jac - j0).max() <= JAC_TOL
        unused = np.arange(K)[None, :] >= cnt[:, None]
        assert not err[unused].any() and not jac[unused].any()
    # the optimiser moves the poses between iterations; the association stays on the device
    poses2 = poses + rng.normal(0, 0.02, poses.shape)
    err, jac = handle.obstacle_edges(poses2, K, ObstacleEdgeParams.make(*info))
    e0, j0 = port.obstacle_edges(poses2, pts, idx, cnt, robot, info)
    assert np.abs(err - e0).max() <= ERR_TOL and np.abs(jac - j0).max() <= JAC_TOL


def test_call_order_and_limits(handle):
    pts = np.random.default_rng(1).uniform(-1, 1, (50, 2))
    poses = np.zeros((4, 3))
    with pytest.raises(PPError) as e:
        handle.obstacle_edges(poses, 5)            # no robot model
    assert e.value.status == 2
    handle.robot_model_set([0.3])
    with pytest.raises(PPError) as e:
        handle.obstacle_edges(poses, 5)            # no association
    assert e.value.status == 2
    handle.radius_topk(pts, poses[:, :2], 0.5, 5)
    with pytest.raises(PPError):
        handle.obstacle_edges(poses, 6)            # K differs from the association's
    with pytest.raises(PPError):
        handle.robot_model_set([0.1, 0.2, 0.3])    # odd count
    with pytest.raises(PPError) as e:
        handle.robot_model_set(np.zeros(2 * 33))   # 33 vertices
    assert e.value.status == 4
    handle.obstacle_edges(poses, 5)
