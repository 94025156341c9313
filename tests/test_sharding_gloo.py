"""CPU, world_size 2 over gloo: the host-side logic of the N > 1 path -- bin ranges, the
max-over-ranks timing reduction and the slab gather -- with the oracle standing in for the kernel.
The gathered volume must be byte-identical to the single-rank volume."""
import os
import socket

import numpy as np
import pytest

from path_planner_b200.sharding import bin_range, twin_chunks


def gather_slabs(local_slab, dist):
    """Stand-in for pp_se2_volume_allgather on host tensors (gloo): every rank contributes its
    [bins_per_rank, H, W] uint8 slab and receives the whole [n_theta, H, W] volume."""
    import torch
    world = dist.get_world_size()
    slab = torch.as_tensor(local_slab).contiguous()
    out = torch.empty((world * slab.shape[0],) + tuple(slab.shape[1:]), dtype=slab.dtype)
    dist.all_gather_into_tensor(out, slab)
    return out


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


def test_twin_chunks_partition_the_bins_and_keep_pairs_together():
    for n in (72, 144, 9, 2):
        for world in (1, 2, 3, 4, 8):
            for chunks in (1, 3, 5):
                seen = []
                for g in range(world):
                    per_rank = twin_chunks(n, g, world, chunks)
                    assert len(per_rank) == chunks
                    for c in per_rank:
                        if n % 2 == 0:  # a bin and its twin sit in the same chunk of the same rank
                            assert all(((b + n // 2) % n) in c for b in c)
                        seen += c
                assert sorted(seen) == list(range(n))
            sizes = [sum(len(c) for c in twin_chunks(n, g, world, 1)) for g in range(world)]
            assert max(sizes) - min(sizes) <= (2 if n % 2 == 0 else 1)


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
    from path_planner_b200.sharding import max_over_ranks
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
        # the pair-sharded, chunked exchange of pp_se2_volume_build_sharded: every chunk of every rank is broadcast
        # from its owner into a full-size volume (gloo broadcasts standing in for the grouped ncclBroadcast calls)
        import torch
        vol = torch.zeros(want.shape, dtype=torch.uint8)
        n_chunks = 3
        for c in range(n_chunks):
            for b in twin_chunks(72, rank, world, n_chunks)[c]:
                vol[b] = torch.from_numpy(p.se2_volume(cost, pts, 72, t0=b, t1=b + 1)[0])
            for root in range(world):
                for b in twin_chunks(72, root, world, n_chunks)[c]:
                    dist.broadcast(vol[b], src=root)
        ok = ok and np.array_equal(vol.numpy(), want)
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
