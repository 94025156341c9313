"""CPU: the obstacle-edge oracle (oracle/teb_edges.c) against the fixture generated from the reference's
own code, and against that code itself where oracle/_ref is built."""
import os

import numpy as np
import pytest

from conftest import ROOT

import oracle

ROBOTS = ("rect", "circle", "pentagon")


@pytest.fixture(scope="module")
def edges_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "teb_edges_v1.npz"))


@pytest.mark.parametrize("name", ROBOTS)
def test_port_matches_reference_fixture(port, edges_golden, name):
    g = edges_golden
    err, jac = port.obstacle_edges(g["poses"], g["pts"], g["idx"], g["cnt"], g[f"robot_{name}"], g["info"])
    assert np.array_equal(err, g[f"err_{name}"])
    assert np.array_equal(jac, g[f"jac_{name}"])
    verts, ri, rc = port.robot_model(g[f"robot_{name}"])
    assert (ri, rc) == tuple(g[f"radii_{name}"])
    # both penalty branches and the inside-the-inscribed-circle branch are exercised
    assert (err[..., 1] == 0).any() and (err[..., 0] > 0).any()


def test_port_matches_reference_on_fresh_inputs(port):
    ref = oracle.load_teb_ref()
    if ref is None:
        pytest.skip("oracle/_ref/libpp_teb_ref.so not built on this machine")
    rng = np.random.default_rng(9)
    pts = rng.uniform(-3, 3, (500, 2))
    poses = np.concatenate([rng.uniform(-2, 2, (30, 2)), rng.uniform(-7, 7, (30, 1))], axis=1)
    idx, d2, cnt, found = port.radius_topk(pts, poses[:, :2], 2.0, 12)
    for robot in ([0.64, 0.46], [0.2], [0.5, 0.0, -0.3, 0.3, -0.3, -0.3], [1.0, 1.0, -1.0, 1.0, -1.0, -1.0, 1.0, -1.0]):
        for info in ([0.1, 0.05, 0.2, 0.1], [0.3, 0.0, 0.6, 0.2]):
            e1, j1 = port.obstacle_edges(poses, pts, idx, cnt, robot, info)
            e2, j2 = ref.obstacle_edges(poses, pts, idx, cnt, robot, info)
            assert np.array_equal(e1, e2) and np.array_equal(j1, j2)
        assert port.robot_model(robot)[1:] == ref.robot_radii(robot)


def test_unused_slots_are_zero_and_bad_models_rejected(port, edges_golden):
    g = edges_golden
    cnt = g["cnt"].copy()
    cnt[::3] = 0
    cnt[1::3] = 5
    err, jac = port.obstacle_edges(g["poses"], g["pts"], g["idx"], cnt, g["robot_rect"], g["info"])
    assert not err[::3].any() and not jac[::3].any() and not err[1::3, 5:].any()
    assert np.array_equal(err[1::3, :5], g["err_rect"][1::3, :5])
    with pytest.raises(ValueError):
        port.robot_model([0.1, 0.2, 0.3])           # odd count (the reference throws)
    with pytest.raises(ValueError):
        port.robot_model([0.1, 0.2, 0.3, 0.4])      # two vertices
