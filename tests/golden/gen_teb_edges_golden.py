"""Generates tests/golden/teb_edges_v1.npz from the reference's OWN obstacle-edge code
(oracle/_ref/libpp_teb_ref.so: robot_model.cpp, commons.cpp, edge_obstacle.hpp compiled unmodified
from /root/reference against the Eigen/g2o stand-ins of oracle/shim_teb). Run in the build container:

    python tests/golden/gen_teb_edges_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from path_planner_b200.synth import make_teb_scene  # noqa: E402

ROBOTS = {
    "rect": [0.72, 0.50],                                                  # path_follower/config/params.yaml:30
    "circle": [0.35],
    "pentagon": [0.4, 0.3, 0.4, -0.3, -0.4, -0.3, -0.5, 0.0, -0.4, 0.3],
}
INFO = [0.1, 0.05, 0.2, 0.1]                                               # commons.h:219-222


def scene(n_points=2000, n_poses=40, seed=3):
    pts, xy = make_teb_scene(n_points, n_poses, seed=seed)
    rng = np.random.default_rng(seed)
    poses = np.concatenate([xy, rng.uniform(-3.3, 3.3, (n_poses, 1))], axis=1)
    return pts, np.ascontiguousarray(poses)


def main():
    ref = oracle.load_teb_ref()
    port = oracle.load_port()
    assert ref is not None, "oracle/_ref/libpp_teb_ref.so is not built (needs /root/reference)"
    pts, poses = scene()
    # a search radius wide enough that some kept obstacles lie beyond both penalty ranges (zero error)
    idx, d2, cnt, found = port.radius_topk(pts, poses[:, :2], 1.6, 16)
    out = {"pts": pts, "poses": poses, "idx": idx, "cnt": cnt, "info": np.array(INFO), "radius_arg": 1.6, "K": 16}
    for name, robot in ROBOTS.items():
        err, jac = ref.obstacle_edges(poses, pts, idx, cnt, robot, INFO)
        out[f"robot_{name}"] = np.array(robot)
        out[f"radii_{name}"] = np.array(ref.robot_radii(robot))
        out[f"err_{name}"] = err
        out[f"jac_{name}"] = jac
        print(name, "radii", out[f"radii_{name}"], "zero-error edges", int((err[..., 1] == 0).sum()), "of", int(cnt.sum()))
    path = os.path.join(ROOT, "tests", "golden", "teb_edges_v1.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
