"""GPU, 2 devices: heading-sharded build + pp_se2_volume_allgather (NCCL), and the fused build that
stores into every rank's volume over peer mappings, must both give every rank a volume
byte-identical to the single-GPU one. Skipped on single-GPU boxes."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from path_planner_b200 import capi
    from path_planner_b200.sharding import bin_range
    from path_planner_b200.synth import make_costmap
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        cost = make_costmap(512, 384, seed=12)
        h = capi.Handle(rank)
        h.costmap_upload(cost)
        h.footprint_from_boundary(0.05, [0.64, 0.46], True)
        obj = [capi.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        h.comm_init(obj[0], rank, world)
        t0, t1 = bin_range(72, rank, world)
        h.se2_volume_build(72, 0.5, 0.5, 0.0, t0, t1)
        h.se2_volume_allgather()
        h.sync()
        gathered = h.se2_volume_download()
        # fused build + exchange: every rank's kernels store into every rank's volume (CUDA IPC peers)
        h2 = capi.Handle(rank)
        h2.costmap_upload(cost)
        h2.footprint_from_boundary(0.05, [0.64, 0.46], True)
        h2.se2_volume_reserve(72)
        handles = [None] * world
        dist.all_gather_object(handles, h2.se2_volume_ipc_export())
        h2.se2_volume_ipc_open(handles, rank, world)
        h2.se2_volume_set_push(True)
        h2.se2_volume_build(72, 0.5, 0.5, 0.0, t0, t1)
        h2.sync()
        dist.barrier()          # every rank's slab has landed everywhere
        pushed = h2.se2_volume_download()
        dist.barrier()          # nobody overwrites its volume while a peer may still read it
        # spatial sharding: every rank builds all 72 bins of its row band and pushes them to everyone
        from path_planner_b200.sharding import row_band
        y0, y1 = row_band(cost.shape[0], rank, world)
        h2.se2_volume_set_rows(y0, y1)
        h2.costmap_upload(cost[::-1].copy())      # a different map, so stale data cannot pass
        h2.se2_volume_build(72)
        h2.sync()
        dist.barrier()
        banded = h2.se2_volume_download()
        dist.barrier()
        h.costmap_upload(cost[::-1].copy())
        h.se2_volume_build(72)
        banded_ok = bool(np.array_equal(banded, h.se2_volume_download()))
        h.costmap_upload(cost)
        h.se2_volume_build(72)  # the whole volume on this GPU alone
        single = h.se2_volume_download()
        # pair-sharded build with the exchange pipelined behind it: NCCL broadcasts, then copy-engine pushes
        sharded_ok = True
        h2.se2_volume_set_push(False)
        h2.se2_volume_set_rows(0, 0)
        h2.costmap_upload(cost)
        obj2 = [capi.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj2, src=0)
        h2.comm_init(obj2[0], rank, world)
        for push in (False, True):
            for chunks in (1, 3):
                h2.costmap_upload(cost[::-1].copy())  # fill the volume with another map's values first
                h2.se2_volume_build(72)
                h2.costmap_upload(cost)
                h2.sync()
                dist.barrier()
                h2.se2_volume_build_sharded(72, chunks, push=push)
                h2.sync()
                sharded_ok = sharded_ok and bool(np.array_equal(h2.se2_volume_download(), single))
                dist.barrier()
        q.put((rank, bool(np.array_equal(gathered, single)) and bool(np.array_equal(pushed, single)) and banded_ok and sharded_ok,
               int(gathered.sum() % 1000003)))
        dist.barrier()
        h2.close()
        h.close()
    finally:
        dist.destroy_process_group()


def test_two_gpu_allgather_is_byte_identical():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res[0][1] and res[1][1]
    assert res[0][2] == res[1][2]
