"""Generates tests/golden/ingest_v1.npz: the reference's own map asset
(src/path_follower/maps/sim.pgm + sim.yaml, read from /root/reference) pushed through the oracle's
ingestion restatement (oracle/ingest.c). Run in the build container only:

    python tests/golden/gen_ingest_golden.py

The image thresholding is cross-checked here against an independent numpy restatement of
map_manager_ros.cpp:93-108. The two costmap_2d layer steps have no reference-side pin (costmap_2d is
not vendored): those arrays pin the oracle against regressions, nothing more.
"""
import os
import re
import sys

import numpy as np
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

MAPS = "/root/reference/src/path_follower/maps"


def read_pgm(path):
    d = open(path, "rb").read()
    m = re.match(rb"P5\s+(?:#[^\n]*\n)?\s*(\d+)\s+(\d+)\s+(\d+)\s", d)
    w, h, maxval = int(m.group(1)), int(m.group(2)), int(m.group(3))
    assert maxval == 255
    return np.frombuffer(d[m.end():], np.uint8).reshape(h, w).copy()


def numpy_threshold(image, occ_th, free_th, negate):
    v = image.astype(np.int64)
    if negate:
        v = 255 - v
    mid = ((v - free_th) / (occ_th - free_th) * 100).astype(np.int64)
    out = np.where(v > occ_th, 100, np.where(v < free_th, 0, mid))
    return out[::-1].astype(np.int8)


def main():
    port = oracle.load_port()
    meta = yaml.safe_load(open(os.path.join(MAPS, "sim.yaml")))
    image = read_pgm(os.path.join(MAPS, meta["image"]))
    occ_th, free_th, negate = float(meta["occupied_thresh"]), float(meta["free_thresh"]), bool(meta["negate"])
    occ = port.occupancy_from_image(image, occ_th, free_th, negate)
    assert np.array_equal(occ, numpy_threshold(image, occ_th, free_th, negate))
    # a second image with every gray level, so the interpolation branch is pinned too
    ramp = (np.add.outer(np.arange(64), np.arange(256)) % 256).astype(np.uint8)
    ramp_occ = port.occupancy_from_image(ramp, occ_th, free_th, False)
    assert np.array_equal(ramp_occ, numpy_threshold(ramp, occ_th, free_th, False))
    lp = oracle.layer_params(resolution=float(meta["resolution"]))
    out = {
        "image": image, "occ_th": occ_th, "free_th": free_th, "negate": negate, "resolution": float(meta["resolution"]),
        "occ": occ, "ramp": ramp, "ramp_occ": ramp_occ,
        "cost_nearest": port.costmap_from_occupancy(occ, lp, "nearest"),
        "cost_propagate": port.costmap_from_occupancy(occ, lp, "propagate"),
    }
    path = os.path.join(ROOT, "tests", "golden", "ingest_v1.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", {k: getattr(v, "shape", v) for k, v in out.items()})
    print("nearest != propagate in", int((out["cost_nearest"] != out["cost_propagate"]).sum()), "cells")


if __name__ == "__main__":
    main()
