"""Generates tests/golden/hotpath_v1.npz from the reference's OWN code (oracle/_ref/libpp_ref.so,
compiled unmodified from /root/reference by oracle/Makefile). Run in the build container only:

    python tests/golden/gen_golden.py

The reference has no golden vectors of its own; these are the pins.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from path_planner_b200.synth import make_costmap, make_teb_scene  # noqa: E402

RES = 0.05
FOOTPRINTS = {
    "rect_hole": ([0.64, 0.46], True),    # shipped default, params.yaml:29-30
    "rect_solid": ([0.64, 0.46], False),
    "box_hole": ([-0.2, -0.25, 0.6, 0.25], True),
    "circle_hole": ([0.35], True),
}
OFFSETS = {"centre": (0.5, 0.5, 0.0), "corner": (0.0, 0.0, 0.0), "frac": (0.37, 0.81, 0.29)}


def main():
    ref = oracle.load_ref()
    assert ref is not None, "oracle/_ref is not built (needs /root/reference)"
    out = {}
    cost = make_costmap(96, 80, seed=11, coverage=0.10, pockets=[(10, 10, 5), (85, 70, 5)])
    out["cost"] = cost
    for fname, (b, hole) in FOOTPRINTS.items():
        out[f"fp_{fname}"] = ref.footprint_points(RES, b, hole)
        for oname, off in OFFSETS.items():
            if fname != "rect_hole" and oname != "centre":
                continue
            out[f"vol_{fname}_{oname}_72"] = ref.se2_volume(cost, RES, b, hole, 72, *off)
    out["vol_rect_hole_centre_144_bins"] = np.arange(0, 144, 7, dtype=np.int32)
    v144 = ref.se2_volume(cost, RES, [0.64, 0.46], True, 144, 0.5, 0.5, 0.0)
    out["vol_rect_hole_centre_144"] = v144[out["vol_rect_hole_centre_144_bins"]]

    rng = np.random.default_rng(5)
    poses = np.stack([rng.uniform(-3, 99, 400), rng.uniform(-3, 83, 400), rng.uniform(-7, 7, 400)], 1).astype(np.float32)
    c, f, k = ref.footprint_cost_batch(cost, RES, [0.64, 0.46], True, poses)
    out["batch_poses"], out["batch_cost"], out["batch_flag"], out["batch_collision"] = poses, c, f, k

    # heuristic: (a) as the planner runs it (stops when the start closes, then lazy queries),
    # (b) forced to full expansion by a start on the border
    start, goal = (10.3, 10.7), (85.2, 70.9)
    q = np.stack([rng.uniform(1, 95, 40), rng.uniform(1, 79, 40)], 1).astype(np.float32)
    ok, table, qo = ref.heuristic(cost, RES, start, goal, 2.0, q)
    out["heur_start"], out["heur_goal"], out["heur_queries"] = np.float32(start), np.float32(goal), q
    out["heur_ok"], out["heur_table"], out["heur_query_out"] = np.int32(ok), table, qo
    ok2, table2, _ = ref.heuristic(cost, RES, (0.0, 0.0), goal, 2.0, None)
    out["heur_full_ok"], out["heur_full_table"] = np.int32(ok2), table2

    pts, qposes = make_teb_scene(2000, 40, seed=3)
    idx, d2, cnt, found = ref.radius_topk(pts, qposes, 0.9574, 10)
    out["teb_pts"], out["teb_queries"] = pts, qposes
    out["teb_idx"], out["teb_d2"], out["teb_cnt"], out["teb_found"] = idx, d2, cnt, found

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hotpath_v1.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
