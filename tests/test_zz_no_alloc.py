"""GPU: pp_heuristic_build_noalloc, the entry point for callers that may not allocate device memory
(real-time planner threads, plugins in a process whose memory pool owns the whole device). It
takes device pointers and keeps its scratch in a fixed array inside the library, so the checks
here are: same field as pp_heuristic_build, bit for bit (the tolerance this suite uses for that
kernel), on two calls in a row that share the scratch -- the second on a different map size and
goal, so stale flags from the first would show -- and a clear refusal beyond the size limit.
The file name sorts last on purpose: a failure here must not hide the rest of the suite."""
import numpy as np
import pytest

from path_planner_b200.synth import make_costmap

pytestmark = pytest.mark.gpu


def run_noalloc(cost, goal, penalty):
    import torch
    from path_planner_b200 import capi
    H, W = cost.shape
    d_cost = torch.from_numpy(cost).cuda()
    d_field = torch.empty((H, W), dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream  # the uploads above are ordered before the call
    capi.heuristic_build_noalloc(d_cost.data_ptr(), W, H, goal[0], goal[1], penalty, d_field.data_ptr(), stream)
    return d_field.cpu().numpy()  # the entry point returns only once the field is complete


def test_noalloc_matches_existing_entry_point_twice_in_a_row(handle, port):
    cases = [((700, 500), (611, 402), 2.0, 21), ((131, 77), (5, 70), 2.1, 22)]
    for (W, H), goal, pen, seed in cases:
        cost = make_costmap(W, H, seed=seed, coverage=0.08, pockets=[(goal[0], goal[1], 3)])
        got = run_noalloc(cost, goal, pen)
        handle.costmap_upload(cost)
        want = handle.heuristic_build(goal[0], goal[1], pen)
        assert np.array_equal(got, want), ((W, H), goal)
        # and not only against the library's other entry point: the CPU checker's converged field
        assert np.array_equal(got, port.heuristic_fixed_point(cost, goal, pen)), ((W, H), goal)
        assert (got > 0).mean() > 0.3
    # and once more on the first case: the scratch now holds the leftovers of a smaller map
    (W, H), goal, pen, seed = cases[0]
    cost = make_costmap(W, H, seed=seed, coverage=0.08, pockets=[(goal[0], goal[1], 3)])
    handle.costmap_upload(cost)
    assert np.array_equal(run_noalloc(cost, goal, pen), handle.heuristic_build(goal[0], goal[1], pen))


def test_noalloc_rejects_oversize_and_bad_arguments():
    import torch
    from path_planner_b200 import capi
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    f = torch.zeros(64, dtype=torch.float32, device="cuda")
    with pytest.raises(capi.PPError) as ei:
        capi.heuristic_build_noalloc(d.data_ptr(), 4097, 4096, 1, 1, 2.0, f.data_ptr(), None)  # 129*128 tiles
    assert ei.value.status == 4  # PP_ERR_RANGE
    with pytest.raises(capi.PPError):
        capi.heuristic_build_noalloc(d.data_ptr(), 8, 8, 9, 0, 2.0, f.data_ptr(), None)  # goal outside
    with pytest.raises(capi.PPError):
        capi.heuristic_build_noalloc(None, 8, 8, 1, 1, 2.0, f.data_ptr(), None)
    cost = np.zeros((32 * 4, 32 * 4), np.uint8)
    assert (run_noalloc(cost, (10, 10), 0.0)[1:-1, 1:-1] > 0).all()


def test_noalloc_at_the_size_limit(handle):
    """4096 x 4096 = 16384 tiles: every slot of the built-in array's key and work-list sections is in use and the
    travel table sits right behind them. The goal is in the last tile, so the wave crosses the whole tile grid."""
    W = H = 4096
    goal = (4090, 4085)
    cost = make_costmap(W, H, seed=5, coverage=0.1, pockets=[(goal[0], goal[1], 4)])
    got = run_noalloc(cost, goal, 2.0)
    handle.costmap_upload(cost)
    want = handle.heuristic_build(goal[0], goal[1], 2.0)
    assert np.array_equal(got, want)
    assert (got[:64, :64] > 0).any()  # the far corner was reached
