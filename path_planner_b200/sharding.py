"""Sharding of the SE2 volume over the GPUs of one box. Two partitions of the same
independent units: heading bins -- rank g of `world` owns bins [g*n/world, (g+1)*n/world), or, with
pp_se2_volume_build_sharded, runs of the pairs (u, u + n/2) (twin_chunks below) -- and map-row
bands (row_band below; what bench.py --gpus N times as `value`). Host-side helpers shared by bench.py and
the tests; the data path itself (pp_se2_volume_build per rank, pp_se2_volume_set_rows,
pp_se2_volume_allgather) is in the library."""


def bin_range(n_theta, rank, world):
    """Half-open range of heading bins of `rank`. The slab all-gather needs equal counts, so
    n_theta must be a multiple of world (72 and 144 are, for 1/2/4/8)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"rank {rank} of {world}")
    if n_theta % world:
        raise ValueError(f"{n_theta} heading bins do not split evenly over {world} ranks")
    per = n_theta // world
    return rank * per, (rank + 1) * per


def max_over_ranks(value, dist=None, device="cpu"):
    """max of a Python float over all ranks (the timing reduction of bench.py)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    import torch
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def twin_chunks(n_theta, rank, world, n_chunks=1):
    """The heading bins of `rank` in pp_se2_volume_build_sharded, chunk by chunk: the n_theta/2 pairs of bins
    (u, u + n_theta/2) are dealt out in contiguous runs (single bins when n_theta is odd), rank g owning pairs
    [g*U/world, (g+1)*U/world); a rank's run is cut into n_chunks parts, the first half as long as the others. Returns a list of lists of bins."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"rank {rank} of {world}")
    pairs = n_theta % 2 == 0
    U = n_theta // 2 if pairs else n_theta
    r0, r1 = U * rank // world, U * (rank + 1) // world
    out = []
    total = 2 * n_chunks - 1  # the first part is half as long as the others (nothing travels while it is built)
    cut = lambda c: 0 if c == 0 else 2 * c - 1
    for c in range(n_chunks):
        a = r0 + ((r1 - r0) * cut(c) + total // 2) // total
        b = r1 if c + 1 == n_chunks else r0 + ((r1 - r0) * cut(c + 1) + total // 2) // total
        bins = []
        for u in range(a, b):
            bins.append(u)
            if pairs:
                bins.append(u + U)
        out.append(bins)
    return out


def row_band(height, rank, world, tile_rows=32):
    """Half-open range of map rows of `rank` for spatial sharding: bands are whole multiples of 32 rows
    (a multiple of the kernel's 16-row tile height, also of the optional 32-row build), as equal as the
    count allows; together they cover [0, height)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"rank {rank} of {world}")
    tiles = (height + tile_rows - 1) // tile_rows
    lo = (tiles * rank) // world * tile_rows
    hi = (tiles * (rank + 1)) // world * tile_rows
    return min(lo, height), min(hi, height)
