"""Sharding of the SE2 volume over the GPUs of one box. Two partitions of the same
independent units: heading bins -- rank g of `world` owns bins [g*n/world, (g+1)*n/world) -- and map-row
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


def gather_slabs(local_slab, dist):
    """Emulation of pp_se2_volume_allgather on host tensors (gloo): every rank contributes its
    [bins_per_rank, H, W] uint8 slab and receives the whole [n_theta, H, W] volume."""
    import torch
    world = dist.get_world_size()
    slab = torch.as_tensor(local_slab).contiguous()
    out = torch.empty((world * slab.shape[0],) + tuple(slab.shape[1:]), dtype=slab.dtype)
    dist.all_gather_into_tensor(out, slab)
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
