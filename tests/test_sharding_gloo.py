"""CPU, world_size 2 over gloo: the host-side logic of the N > 1 path -- bin ranges, the
max-over-ranks timing reduction and the slab gather -- with the oracle standing in for the kernel.
The gathered volume must be byte-identical to the single-rank volume."""
import os
import socket

import numpy as np
import pytest

from path_planner_b200.sharding import bin_range


def test_bin_ranges_cover_exactly():
    for n in (72, 144):
        for world in (1, 2, 4, 8):
            r = [bin_range(n, g, world) for g in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))
            assert len({b - a for a, b in r}) == 1
    with pytest.raises(ValueError):
        bin_range(72, 0, 5)
    with pytest.raises(ValueError):
        bin_range(72, 2, 2)


def test_row_bands_cover_exactly():
    from path_planner_b200.sharding import row_band
    for H in (1, 31, 32, 33, 400, 4096, 8192):
        for world in (1, 2, 4, 8):
            r = [row_band(H, g, world) for g in range(world)]
            assert r[0][0] == 0 and r[-1][1] == H
            assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))
            assert all(a % 32 == 0 for a, _ in r)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch.distributed as dist
    import oracle
    from path_planner_b200.sharding import gather_slabs, max_over_ranks
    from path_planner_b200.synth import make_costmap
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cost = make_costmap(96, 64, seed=5)
        p = oracle.load_port()
        pts = p.footprint_points(0.05, [0.64, 0.46], True)
        t0, t1 = bin_range(72, rank, world)
        slab = p.se2_volume(cost, pts, 72, t0=t0, t1=t1)
        full = gather_slabs(slab, dist).numpy()
        want = p.se2_volume(cost, pts, 72)
        ok = np.array_equal(full, want)
        slowest = max_over_ranks(1.0 + rank, dist)
        q.put((rank, bool(ok), slowest))
    finally:
        dist.destroy_process_group()


def test_two_rank_gather_is_byte_identical():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [r[1] for r in res] == [True, True]
    assert [r[2] for r in res] == [2.0, 2.0]  # both ranks see the slowest rank's time
