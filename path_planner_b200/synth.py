"""Seeded synthetic inputs for tests and bench.py. Harness-side only (numpy);
nothing here is on the product's compute path."""
import numpy as np


def make_costmap(W, H, seed=0, coverage=0.12, unknown_frac=0.005, decay=0.4, inflate_cells=10, pockets=()):
    """uint8[H][W] costmap_2d-style grid: 0 free, 254 lethal rectangles/discs covering ~`coverage`,
    253 in the one-cell ring around them, max(1, 252*exp(-decay*(d-1))) out to `inflate_cells`,
    `unknown_frac` of cells 255 (NO_INFORMATION). `pockets` = [(x, y, r), ...] are cleared to 0."""
    from scipy import ndimage

    rng = np.random.default_rng(seed)
    lethal = np.zeros((H, W), bool)
    target = coverage * W * H
    area = 0.0
    smax = max(4, min(48, min(W, H) // 8))
    while area < target:
        w = int(rng.integers(2, smax))
        h = int(rng.integers(2, smax))
        x = int(rng.integers(0, max(1, W - w)))
        y = int(rng.integers(0, max(1, H - h)))
        if rng.random() < 0.3:
            yy, xx = np.ogrid[:h, :w]
            disc = (xx - (w - 1) / 2) ** 2 / max(1.0, (w / 2) ** 2) + (yy - (h - 1) / 2) ** 2 / max(1.0, (h / 2) ** 2) <= 1.0
            lethal[y:y + h, x:x + w] |= disc[: H - y, : W - x]
        else:
            lethal[y:y + h, x:x + w] = True
        area += 0.8 * w * h
    d = ndimage.distance_transform_edt(~lethal)
    cost = np.zeros((H, W), np.uint8)
    infl = (d > 0) & (d <= inflate_cells)
    val = np.maximum(1.0, np.rint(252.0 * np.exp(-decay * (d - 1.0))))
    cost[infl] = np.minimum(252, val[infl]).astype(np.uint8)
    cost[(d > 0) & (d < 1.5)] = 253
    cost[lethal] = 254
    unk = rng.random((H, W)) < unknown_frac
    cost[unk] = 255
    for (px, py, r) in pockets:
        yy, xx = np.ogrid[:H, :W]
        cost[(xx - px) ** 2 + (yy - py) ** 2 <= r * r] = 0
    return np.ascontiguousarray(cost)


def make_teb_scene(n_points=10000, n_poses=100, seed=0, max_follow_range=4.5):
    """N obstacle points uniform in the +-1.2*max_follow_range box and Q poses along a gently
    curving 8 m path."""
    rng = np.random.default_rng(seed)
    half = 1.2 * max_follow_range
    pts = rng.uniform(-half, half, size=(n_points, 2))
    s = np.linspace(0.0, 8.0, n_poses)
    poses = np.stack([s - 4.0, 0.8 * np.sin(s * 0.6)], axis=1)
    return np.ascontiguousarray(pts), np.ascontiguousarray(poses)
