#!/usr/bin/env python
"""bench.py -- the reference's headline workload on B200: SE2 collision volume (+ heuristic field,
TEB queries) at 4096x4096x72 (BASELINE.json configs[2]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One process per GPU (torchrun for N > 1). A step is one build of the whole SE2 volume from a costmap
already resident in HBM; with N > 1 every rank builds all 72 heading bins of its band of map rows
(strong scaling, independent units, no data-path collective in `value`; heading-bin sharding, its
NCCL all-gather and the fused peer-store variants are timed separately and reported under
"gathered"). Rank 0 prints ONE JSON line.

`--impl reference` times the reference's own CPU code (oracle/_ref, compiled unmodified from the
reference sources; falls back to the oracle port where that is absent) with all host threads on
a bounded row band of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W = H = 4096
N_THETA = 72
RES = 0.05
BOUNDARY, HOLE = [0.64, 0.46], True  # shipped default footprint, path_planner/config/params.yaml:29-30
COST_PENALTY = 2.0
GOAL = (3000, 1000)
METRIC = "se2_collision_cells_per_sec"
UNIT = "cells/s"
WORKLOAD = "se2_volume_4096x4096x72_rect0.64x0.46_hole"


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: an NVML polling thread (5 ms
    period, so even a short region gets samples), falling back to `nvidia-smi -lms` where NVML is
    not importable."""
    REASONS = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40, "sw_power_cap": 0x4}

    def __init__(self, gpu_index):
        import threading
        self.sm, self.reason_bits, self.max_mhz = [], 0, None
        self.stop_flag = False
        self.thread = None
        self.proc = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.dev = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.dev, pynvml.NVML_CLOCK_SM))
            sys.setswitchinterval(5e-4)  # the timed loop is a tight run of short ctypes calls: let the poller in often
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
        except Exception:
            self.nv = None
            self.tmp = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
            try:
                self.proc = subprocess.Popen(
                    ["nvidia-smi", "-i", str(gpu_index), "--query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.active",
                     "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.tmp, stderr=subprocess.DEVNULL)
            except Exception:
                self.proc = None

    def _poll(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.dev, nv.NVML_CLOCK_SM)))
                self.reason_bits |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.dev))
            except Exception:
                pass
            time.sleep(0.005)

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.thread is not None:
            self.stop_flag = True
            self.thread.join(timeout=2)
            if self.sm:
                out["sm_mhz"] = float(np.median(self.sm))
                out["sm_max_mhz"] = self.max_mhz
                out["samples"] = len(self.sm)
            out["reasons"] = sorted(k for k, bit in self.REASONS.items() if self.reason_bits & bit)
            out["source"] = "nvml"
            return out
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.tmp.flush()
        self.tmp.seek(0)
        sm, mx, bits = [], [], 0
        for line in self.tmp.read().splitlines():
            p = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(p[0]))
                mx.append(float(p[1]))
                bits |= int(p[2], 16)
            except (ValueError, IndexError):
                continue
        try:
            os.unlink(self.tmp.name)
        except OSError:
            pass
        if sm:
            out["sm_mhz"], out["sm_max_mhz"], out["samples"] = float(np.median(sm)), float(max(mx)), len(sm)
        out["reasons"] = sorted(k for k, bit in self.REASONS.items() if bits & bit)
        out["source"] = "nvidia-smi"
        return out


def make_inputs():
    from path_planner_b200.synth import make_costmap
    return make_costmap(W, H, seed=1, pockets=[(GOAL[0], GOAL[1], 6)])


PLAN_SCENES = {
    # BASELINE.json configs[0] and configs[1]: Hybrid A* on 400x400 (rect footprint) and 2048x2048 (box footprint)
    "config1_400x400_rect": dict(W=400, seed=3, coverage=0.06, start=(2.0, 2.0, 0.3), goal=(17.5, 17.0, 1.2),
                                 boundary=[0.64, 0.46], params={"motion_type": 2}),
    "config2_2048x2048_box": dict(W=2048, seed=2, coverage=0.04, start=(5.0, 5.0, 0.5), goal=(60.0, 45.0, 0.0),
                                  boundary=[-0.2, -0.25, 0.6, 0.25], params={"motion_type": 2, "max_iterations": 400000}),
}


def occupancy_of(cost):
    """nav_msgs/OccupancyGrid data whose static layer has the bench costmap's obstacles and unknown cells."""
    occ = np.zeros(cost.shape, np.int8)
    occ[cost == 254] = 100
    occ[cost == 255] = -1
    return occ


def plan_scene(sc):
    from path_planner_b200.synth import make_costmap
    pockets = [(int(sc["start"][0] / RES), int(sc["start"][1] / RES), 25), (int(sc["goal"][0] / RES), int(sc["goal"][1] / RES), 25)]
    return make_costmap(sc["W"], sc["W"], seed=sc["seed"], coverage=sc["coverage"], unknown_frac=0.0, pockets=pockets)


def cpu_volume_sample(cost, target_seconds, threads):
    """Reference CPU volume on a row band sized for ~target_seconds; returns (cells/s, kind, sample)."""
    import oracle
    ref = oracle.load_ref()
    y0 = 1024
    if ref is not None:
        kind = "reference"
        run = lambda rows: ref.se2_volume(cost, RES, BOUNDARY, HOLE, N_THETA, y0=y0, y1=y0 + rows, nthreads=threads)
    else:
        kind = "port"
        port = oracle.load_port()
        pts = port.footprint_points(RES, BOUNDARY, HOLE)
        run = lambda rows: port.se2_volume(cost, pts, N_THETA, y0=y0, y1=y0 + rows, nthreads=threads)
    t = time.perf_counter()
    run(4)
    probe = time.perf_counter() - t
    rows = int(max(4, min(2048, 4 * target_seconds / max(probe, 1e-3))))
    t = time.perf_counter()
    run(rows)
    dt = time.perf_counter() - t
    cells = rows * W * N_THETA
    return cells / dt, kind, f"rows [{y0},{y0 + rows}) of {H} x all {W} columns x all {N_THETA} bins ({cells} cells, {dt:.1f} s)"


def run_reference(args, rank):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    cost = make_inputs()
    per_step = max(0.5, min(20.0, 90.0 / max(1, args.steps + args.warmup)))  # the whole arm stays within a few minutes
    vals = []
    kind = sample = None
    for i in range(args.warmup + args.steps):
        v, kind, sample = cpu_volume_sample(cost, per_step, cores)
        if i >= args.warmup:
            vals.append(v)
    value = float(np.mean(vals))
    total_cells = W * H * N_THETA
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total_cells / value, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "CPU; each step is a row band, ms_per_step extrapolated to the full volume"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--mode", type=int, default=0, help="0 auto, 1 force the generic kernel")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    if N_THETA % world:
        raise SystemExit(f"{N_THETA} heading bins do not split evenly over {world} ranks")

    # Libraries (NCCL prints its version on stdout) must not break the one-JSON-line contract: send
    # everything that lands on fd 1 to stderr and keep the real stdout for the result line.
    sys.stdout.flush()
    result_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    from path_planner_b200 import capi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    cost = make_inputs()
    from path_planner_b200.sharding import bin_range
    t0, t1 = bin_range(N_THETA, rank, world)
    per = t1 - t0
    h = capi.Handle(local_rank)
    pinned_cost = capi.PinnedBuffer(W * H)
    pinned_cost.array.reshape(H, W)[:] = cost
    cost_pinned = pinned_cost.array.reshape(H, W)
    h.costmap_upload(cost_pinned, pinned=True)
    h.footprint_from_boundary(RES, BOUNDARY, HOLE)
    h.se2_volume_set_mode(args.mode)
    h.sync()
    if world > 1:
        obj = [capi.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        h.comm_init(obj[0], rank, world)

    # N > 1: every rank builds ALL heading bins of its band of map rows -- independent units, no
    # exchange on the data path, and the per-tile table build is paid once per tile in the whole job
    # (heading-bin sharding, SURVEY.md section 8e, repeats it on every rank and separates the bin pairs
    # t / t + n/2 that share one evaluation; it is measured below for comparison).
    from path_planner_b200.sharding import row_band
    y0, y1 = row_band(H, rank, world) if world > 1 else (0, H)
    if world > 1:
        h.se2_volume_set_rows(y0, y1)

    def step():
        h.se2_volume_build(N_THETA, 0.5, 0.5, 0.0, 0, N_THETA)

    def timed(fn, steps):
        """CUDA-event time per step of fn() on this rank's stream, L2 flushed before every step; max over ranks."""
        barrier()
        ms = 0.0
        for _ in range(steps):
            h.l2_flush()
            h.timer_start()
            fn()
            ms += h.timer_stop()
        barrier()
        return max_over_ranks(ms / steps), ms / steps

    # ---- device-resident throughput (`value`) -------------------------------------------------
    for _ in range(args.warmup):
        h.l2_flush()
        step()
    h.sync()
    stats = h.se2_volume_stats()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    launches0 = h.launch_count()
    wall0 = time.perf_counter()
    ms_per_step, my_ms = timed(step, args.steps)
    wall = time.perf_counter() - wall0
    launches = h.launch_count() - launches0
    clocks = sampler.stop() if sampler else None
    total_cells = W * H * N_THETA
    value = total_cells / (ms_per_step * 1e-3)

    # ---- N > 1: the other ways to split and exchange, for comparison ---------------------------------
    gathered = None
    if world > 1:
        def cells_per_s(ms):
            return total_cells / (ms * 1e-3)

        def theta_step():
            h.se2_volume_build(N_THETA, 0.5, 0.5, 0.0, t0, t1)

        def theta_gather_step():
            theta_step()
            h.se2_volume_allgather()

        gathered = {}
        # the row bands again, every output word also stored to all other ranks' volumes (NVLink peer
        # mappings through CUDA IPC): the whole volume on every GPU without a separate collective
        h.se2_volume_reserve(N_THETA)
        handles = [None] * world
        dist.all_gather_object(handles, h.se2_volume_ipc_export())
        h.se2_volume_ipc_open(handles, rank, world)
        h.se2_volume_set_push(True)
        for _ in range(2):
            step()
        h.sync()
        ms, _ = timed(step, args.steps)
        gathered["row_sharded_fused_push"] = {"value": cells_per_s(ms), "unit": UNIT, "ms_per_step": ms,
                                              "how": "row bands, every output word also stored to all other ranks' volumes"}
        h.se2_volume_set_push(False)
        # heading-bin sharding: rank g builds bins [g*n/N, (g+1)*n/N) of the whole map
        h.se2_volume_set_rows(0, 0)
        for _ in range(4):  # a handle re-plans thoroughly on its third build of a configuration: keep that out of the timing
            theta_gather_step()
        h.sync()
        ms, _ = timed(theta_step, args.steps)
        gathered["theta_sharded"] = {"value": cells_per_s(ms), "unit": UNIT, "ms_per_step": ms,
                                     "how": "rank g builds its slab of heading bins of the whole map (no exchange)"}
        ms, _ = timed(theta_gather_step, args.steps)
        gathered["theta_sharded_allgather"] = {"value": cells_per_s(ms), "unit": UNIT, "ms_per_step": ms,
                                               "collective": "ncclAllGather of heading slabs, in place",
                                               "bytes_received_per_gpu": int(W * H * (N_THETA - per))}
        h.se2_volume_set_push(True)
        for _ in range(2):
            theta_step()
        h.sync()
        ms, _ = timed(theta_step, args.steps)
        gathered["theta_sharded_fused_push"] = {"value": cells_per_s(ms), "unit": UNIT, "ms_per_step": ms,
                                                "how": "slab build kernels store every output word to all ranks' volumes; no separate collective"}
        h.se2_volume_set_push(False)
        # pairs of bins (u, u + n/2) dealt out in runs, the exchange of part k hidden behind the build of part
        # k + 1 (pp_se2_volume_build_sharded); next to it the exchange alone on the same volume, so that what is
        # reported is the cost of the link and how much of the build hides behind it
        import zlib
        recv_bytes = int(W * H * (N_THETA - per))
        for key, push in (("theta_pairs_pipelined_nccl", False), ("theta_pairs_pipelined_push", True)):
            best = None
            for chunks in (1, 2, 3, 4):
                def sharded_step():
                    h.se2_volume_build_sharded(N_THETA, chunks, push=push)
                for _ in range(3):
                    sharded_step()
                h.sync()
                ms, _ = timed(sharded_step, args.steps)
                if best is None or ms < best[0]:
                    best = (ms, chunks)
            chunks = best[1]

            def exchange_only():
                h.se2_volume_build_sharded(N_THETA, chunks, push=push, exchange_only=True)
            ms_x, _ = timed(exchange_only, args.steps)
            h.se2_volume_build_sharded(N_THETA, chunks, push=push)
            h.sync()
            gathered[key] = {"value": cells_per_s(best[0]), "unit": UNIT, "ms_per_step": best[0], "chunks": chunks,
                             "exchange_alone_ms": ms_x, "ratio_to_exchange_alone": best[0] / ms_x,
                             "bytes_received_per_gpu": recv_bytes, "inbound_GBps": recv_bytes / (best[0] * 1e-3) / 1e9,
                             "how": ("copy engines into the peers' mapped volumes + one-word all-reduce" if push else
                                     "one group of ncclBroadcast calls per part (2 per owner)")}
        # plain ncclAllGather of equal slabs on an already built volume, for reference
        h.se2_volume_build(N_THETA, 0.5, 0.5, 0.0, t0, t1)
        ms, _ = timed(h.se2_volume_allgather, args.steps)
        gathered["allgather_alone"] = {"ms_per_step": ms, "inbound_GBps": recv_bytes / (ms * 1e-3) / 1e9,
                                       "collective": "ncclAllGather of heading slabs, in place, nothing built"}
        # every rank checks what it now holds against a volume built on ONE GPU (rank 0's CRC-32 per heading bin):
        # the driver's GPU test box has a single GPU, so this is where the multi-GPU bytes are verified
        h.se2_volume_build_sharded(N_THETA, 3, push=False)
        h.sync()
        mine = [zlib.crc32(h.se2_volume_download(t, t + 1).tobytes()) for t in range(N_THETA)]
        h.se2_volume_build(N_THETA)
        h.sync()
        single = [zlib.crc32(h.se2_volume_download(t, t + 1).tobytes()) for t in range(N_THETA)]
        ref_crc = [single if rank == 0 else None]
        dist.broadcast_object_list(ref_crc, src=0)
        ok = torch.tensor([int(mine == ref_crc[0] and single == ref_crc[0])], device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        gathered["verified"] = {"sharded_volume_equals_single_gpu_build_on_every_rank": bool(ok.item()),
                                "how": "CRC-32 per heading bin against rank 0's single-GPU build"}
        h.se2_volume_set_rows(y0, y1)
        for _ in range(4):  # back to the primary configuration (its plan may have been replaced above)
            step()
        h.sync()
        # the row bands are verified the same way: this rank's band of every bin against rank 0's whole volume
        band_crc = [zlib.crc32(np.ascontiguousarray(h.se2_volume_download(t, t + 1)[0, y0:y1]).tobytes()) for t in range(N_THETA)]
        all_crc = [None] * world
        dist.all_gather_object(all_crc, (y0, y1, band_crc))
        if rank == 0:
            h.se2_volume_set_rows(0, 0)
            h.se2_volume_build(N_THETA)
            h.sync()
            good = True
            for t in range(N_THETA):
                plane = h.se2_volume_download(t, t + 1)[0]
                for (a, b, crcs) in all_crc:
                    good = good and zlib.crc32(np.ascontiguousarray(plane[a:b]).tobytes()) == crcs[t]
            gathered["verified"]["row_bands_equal_single_gpu_build"] = bool(good)
            h.se2_volume_set_rows(y0, y1)
            for _ in range(4):
                step()
            h.sync()
        barrier()

    # ---- end to end through the C ABI with host buffers (`e2e`) ----------------------------------
    band = y1 - y0
    out_pinned = capi.PinnedBuffer(W * band * N_THETA)
    out_view = out_pinned.array.reshape(N_THETA, band, W)
    e2e_steps = max(3, min(args.steps, 5))

    def e2e_step():
        h.costmap_upload(cost_pinned, pinned=True)
        step()
        if world > 1:
            h.se2_volume_download_rows(y0, y1, 0, N_THETA, out=out_view)
        else:
            h.se2_volume_download(0, N_THETA, out=out_view, pinned=True)

    e2e_step()
    barrier()
    e0 = time.perf_counter()
    h.timer_start()
    for _ in range(e2e_steps):
        e2e_step()
    e_ms = h.timer_stop()
    barrier()
    e_wall = (time.perf_counter() - e0) * 1e3
    e_ms = max_over_ranks(max(e_ms, 0.0) / e2e_steps)
    e2e = {"value": total_cells / (e_ms * 1e-3), "unit": UNIT, "ms_per_step": e_ms, "wall_ms_per_step": e_wall / e2e_steps,
           "h2d_bytes_per_step": int(W * H * world), "d2h_bytes_per_step": int(W * H * N_THETA),
           "path": "pp_costmap_upload(pinned host) + pp_se2_volume_build + pp_se2_volume_download"
                   + ("_rows(this rank's band; pinned host)" if world > 1 else "(pinned host)")}

    # ---- roofline of the dominant kernel ---------------------------------------------------------
    peak, peak_src = measured_peak_gbs()
    alg_bytes = W * (y1 - y0) * (1 + N_THETA)   # read the costmap once, write every output byte once (this rank's rows)
    achieved = alg_bytes / (my_ms * 1e-3) / 1e9
    traffic = traffic_src = None
    tpath = os.path.join(ROOT, "profiles", "se2_traffic.json")  # written by tools/make_traffic_json.py from an ncu capture
    if os.path.exists(tpath):
        try:
            tj = json.load(open(tpath))
            traffic, traffic_src = tj.get("dram_bytes_per_launch"), tj.get("source")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "kernel": "se2 volume build (all launches of one pp_se2_volume_build)",
                "algorithmic_bytes_per_launch": alg_bytes, "ms_per_launch": my_ms,
                "frac_of_nominal_8TBs": achieved / 8000.0}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "W": W, "H": H, "n_theta": N_THETA, "rows_per_gpu": int(y1 - y0),
                   "footprint_points": int(len(h.footprint_get())), "sample_offset": [0.5, 0.5, 0.0],
                   "sharding": f"row bands x{world}", "l2": "flushed between steps (256 MiB memset, outside the timed events)",
                   "kernel_mode": "auto" if args.mode == 0 else "generic", "fast_tiles": stats["fast_tiles"],
                   "generic_tiles": stats["generic_tiles"], "samples_per_cell": stats["samples_per_cell"]},
        "roofline": roofline, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
        "wall_s_timed_region": wall,
    }
    if gathered:
        line["gathered"] = gathered
    if rank == 0 and world == 1:
        # the same build with the other footprint kinds (BASELINE config 2 uses the asymmetric box: no twin bins)
        line["footprints"] = {}
        hf = capi.Handle(local_rank)
        hf.costmap_upload(cost_pinned, pinned=True)
        for name, (bnd, hole) in {"box_-0.2_-0.25_0.6_0.25_rim": ([-0.2, -0.25, 0.6, 0.25], True), "rect_0.64x0.46_solid": ([0.64, 0.46], False),
                                  "circle_0.35_rim": ([0.35], True)}.items():
            hf.footprint_from_boundary(RES, bnd, hole)
            for _ in range(4):
                hf.se2_volume_build(N_THETA)
            hf.sync()
            ms = 0.0
            for _ in range(10):
                hf.l2_flush()
                hf.timer_start()
                hf.se2_volume_build(N_THETA)
                ms += hf.timer_stop() / 10
            st = hf.se2_volume_stats()
            line["footprints"][name] = {"ms_per_build": ms, "roofline_frac": (W * H * (1 + N_THETA)) / (ms * 1e-3) / 1e9 / peak,
                                        "samples_per_cell": st["samples_per_cell"], "generic_tiles": st["generic_tiles"]}
        hf.close()

    if rank == 0:
        # heuristic field and TEB queries stay on one GPU
        for _ in range(2):
            h.heuristic_build(GOAL[0], GOAL[1], COST_PENALTY, download=False)
        h.sync()
        reps = 5
        h.timer_start()
        for _ in range(reps):
            h.heuristic_build(GOAL[0], GOAL[1], COST_PENALTY, download=False)
        hms = h.timer_stop() / reps
        hs = h.heuristic_stats()
        line["heuristic"] = {"build_ms": hms, "rounds": hs["rounds"], "map": f"{W}x{H}", "unit": "ms"}
        # plan ms: the library's own host Hybrid-A* over GPU tables (pp_plan), second call (buffers warm)
        line["plan"] = {}
        hp = capi.Handle(local_rank)
        for name, sc in PLAN_SCENES.items():
            pc = plan_scene(sc)
            hp.costmap_upload(pc)
            hp.footprint_from_boundary(RES, sc["boundary"], True)
            hp.plan(RES, (0.0, 0.0), sc["start"], sc["goal"], sc["params"])
            r = hp.plan(RES, (0.0, 0.0), sc["start"], sc["goal"], sc["params"])
            line["plan"][name] = {"ms": r["tables_ms"] + r["search_ms"], "tables_ms": r["tables_ms"], "search_ms": r["search_ms"],
                                  "iterations": r["iterations"], "path_poses": int(len(r["path"])), "found": bool(r["ok"])}
            # the same plan with the analytic expansion decided at continuous poses and the path validated exactly,
            # and how often the volume lookup (cell centre, nearest bin) disagrees with the continuous-pose check
            for _ in range(2):  # (a handle re-plans its volume build thoroughly on the third use of a configuration: not timed)
                hp.plan(RES, (0.0, 0.0), sc["start"], sc["goal"], dict(sc["params"], exact_analytic=True, validate_path=True))
            hp.plan_set_query_log(True)
            rx = hp.plan(RES, (0.0, 0.0), sc["start"], sc["goal"], dict(sc["params"], exact_analytic=True, validate_path=True))
            qp, qk, qf, qc = hp.plan_query_log()
            hp.plan_set_query_log(False)
            xc_, xf_ = hp.footprint_cost_batch(qp)
            prim = qk == 0
            line["plan"][name]["exact_analytic"] = {"ms": rx["tables_ms"] + rx["search_ms"], "iterations": rx["iterations"],
                                                    "path_poses_in_collision": rx["stats"]["path_poses_in_collision"],
                                                    "path_poses_checked": rx["stats"]["path_poses_checked"],
                                                    "confirm_batches": rx["stats"]["analytic_confirm_batches"]}
            line["plan"][name]["quantisation"] = {"questions": int(len(qp)), "primitive_false_free": float(((qf == 0) & (xf_ == 1) & prim).mean()),
                                                  "primitive_false_collision": float(((qf == 1) & (xf_ == 0) & prim).mean()),
                                                  "cost_differs": float((qc != xc_.astype(np.uint8)).mean())}
        hp.close()
        # K0: costmap upload over the host link: pinned, pageable (staged), one dirty rectangle
        hu = capi.Handle(local_rank)
        pageable = np.ascontiguousarray(cost)
        up = {}
        for key, src, pinned in (("h2d_pinned_GBps", cost_pinned, True), ("h2d_pageable_staged_GBps", pageable, False)):
            hu.costmap_upload(src, pinned=pinned)
            hu.sync()
            t = time.perf_counter()
            for _ in range(10):
                hu.costmap_upload(src, pinned=pinned)
            hu.sync()
            up[key] = W * H * 10 / (time.perf_counter() - t) / 1e9
        patch = np.ascontiguousarray(cost[1000:1256, 2000:2256])
        hu.costmap_update_rect(2000, 1000, patch)
        hu.sync()
        t = time.perf_counter()
        for _ in range(50):
            hu.costmap_update_rect(2000, 1000, patch)
        hu.sync()
        up["rect_256x256_us"] = 1e6 * (time.perf_counter() - t) / 50
        up["d2h_pinned_GBps_volume_download"] = W * H * N_THETA / (e2e["ms_per_step"] * 1e-3) / 1e9 if world == 1 else None
        up["bytes"] = W * H
        line["upload"] = up
        hu.close()
        # N2 ingestion on the same map: occupancy grid -> static layer + inflation -> costmap, on its own handle
        hi = capi.Handle(local_rank)
        occ = occupancy_of(cost)
        lps = [capi.LayerParams.make(resolution=RES), capi.LayerParams.make(resolution=RES, cost_scaling_factor=8.0 + 1e-9)]
        for k in range(3):
            hi.occupancy_upload(occ)
            hi.costmap_from_occupancy(lps[k % 2])
        hi.sync()
        t = time.perf_counter()
        for k in range(5):
            hi.occupancy_upload(occ)
            hi.costmap_from_occupancy(lps[0])
        hi.sync()
        ing_e2e = 1e3 * (time.perf_counter() - t) / 5
        hi.timer_start()
        for k in range(10):  # a changed parameter forces the full recompute without a new upload
            hi.costmap_from_occupancy(lps[(k + 1) % 2])
        ing_ms = hi.timer_stop() / 10
        patch = np.ascontiguousarray(occ[1000:1256, 2000:2256])
        hi.costmap_from_occupancy(lps[0])
        hi.sync()
        t = time.perf_counter()
        for k in range(20):
            hi.occupancy_update_rect(2000, 1000, patch)
            hi.costmap_from_occupancy(lps[0])
        hi.sync()
        rect_ms = 1e3 * (time.perf_counter() - t) / 20
        # the follower's rolling window (10 m x 10 m at 0.05 m) cut from that map at a moving origin, on its own handle
        hw = capi.Handle(local_rank)
        for k in range(3):
            hw.costmap_window_from_occupancy(hi, 1000 + k, 900, 200, 200, lps[0])
        hw.sync()
        hw.timer_start()
        for k in range(50):
            hw.costmap_window_from_occupancy(hi, 1000 + 7 * k, 900 + 3 * k, 200, 200, lps[0])
        win_ms = hw.timer_stop() / 50
        hw.close()
        line["ingest"] = {"full_recompute_ms": ing_ms, "e2e_ms_with_upload": ing_e2e, "rolling_window_200x200_ms": win_ms,
                          "rect_256x256_e2e_ms": rect_ms, "map": f"{W}x{H}",
                          "inflation_cells": 10, "algorithmic_bytes": 2 * W * H,
                          "note": "full_recompute_ms includes the 101-byte cost-table upload that a parameter change triggers"}
        hi.close()
        from path_planner_b200.synth import make_teb_scene, make_costmap
        pts, poses = make_teb_scene(10000, 100, seed=0)
        h.radius_topk(pts, poses, 0.9574, 10)
        t = time.perf_counter()
        rounds = 30
        for _ in range(rounds):
            h.radius_topk(None, poses, 0.9574, 10)
        line["teb"] = {"us_per_round_e2e": 1e6 * (time.perf_counter() - t) / rounds,
                       "workload": "100 poses x 10000 points, K=10, host buffers in and out"}
        # K3b over larger follower windows: candidates extracted from a costmap stay column-ordered on the device and a
        # query scans only the columns in reach; `full_scan` is the same kernel without the pruning
        sweep = {}
        hs_ = capi.Handle(local_rank)
        for target, (ww, cov) in {"1e4": (500, 0.035), "1e5": (1000, 0.09), "1e6": (2400, 0.16)}.items():
            cm = make_costmap(ww, ww, seed=11, coverage=cov, unknown_frac=0.0)
            hs_.costmap_upload(cm)
            centre = ww * RES / 2
            n_pts = hs_.obstacles_extract(RES, (0.0, 0.0), (centre, centre, 0.0), 1e6, 0.0, download=False)
            q = np.stack([np.linspace(0.1 * ww * RES, 0.9 * ww * RES, 100), centre + 2.0 * np.sin(np.linspace(0, 6, 100))], 1)
            res_ = {}
            for key, env in (("pruned_us", None), ("full_scan_us", "1")):
                if env:
                    os.environ["PP_TEB_NO_PRUNE"] = env
                hs_.radius_topk(None, q, 0.9574, 10)
                t = time.perf_counter()
                for _ in range(10):
                    hs_.radius_topk(None, q, 0.9574, 10)
                res_[key] = 1e6 * (time.perf_counter() - t) / 10
                os.environ.pop("PP_TEB_NO_PRUNE", None)
            sweep[target] = dict(points=int(n_pts), window_cells=ww, **res_)
        hs_.close()
        line["teb"]["n_sweep_100_queries"] = sweep
        # N3: errors + numeric Jacobians of all obstacle edges of one optimiser iteration (association on the device)
        poses3 = np.ascontiguousarray(np.concatenate([poses, 0.6 * np.cos(poses[:, :1])], axis=1))
        h.robot_model_set([0.72, 0.50])
        h.obstacle_edges(poses3, 10)
        t = time.perf_counter()
        for _ in range(rounds):
            h.obstacle_edges(poses3, 10)
        line["teb"]["edges_us_per_iteration_e2e"] = 1e6 * (time.perf_counter() - t) / rounds

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        v, kind, sample = cpu_volume_sample(cost, 15.0, cores)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample}
        try:
            import oracle
            ref = oracle.load_ref()
            if ref is not None:
                ref.heuristic(cost, RES, (0.0, 0.0), (GOAL[0] + 0.5, GOAL[1] + 0.5), COST_PENALTY)
                line["cpu_baseline"]["heuristic_ms"] = 1e3 * ref.last_seconds
                from path_planner_b200.synth import make_teb_scene
                pts, poses = make_teb_scene(10000, 100, seed=0)
                ref.radius_topk(pts, poses, 0.9574, 10)
                line["cpu_baseline"]["teb_us_per_round"] = 1e6 * (ref.last_build_seconds + ref.last_query_seconds)
                teb_ref = oracle.load_teb_ref()
                if teb_ref is not None:  # the reference's own EdgeObstacle + g2o-style numeric Jacobian, 1 thread
                    idx, d2, cnt, found = ref.radius_topk(pts, poses, 0.9574, 10)
                    poses3 = np.ascontiguousarray(np.concatenate([poses, 0.6 * np.cos(poses[:, :1])], axis=1))
                    t = time.perf_counter()
                    for _ in range(20):
                        teb_ref.obstacle_edges(poses3, pts, idx, cnt, [0.72, 0.50], [0.1, 0.05, 0.2, 0.1])
                    line["cpu_baseline"]["teb_edges_us_per_iteration"] = 1e6 * (time.perf_counter() - t) / 20
                port = oracle.load_port()
                olp = oracle.layer_params(resolution=RES)
                t = time.perf_counter()
                port.costmap_from_occupancy(occupancy_of(cost), olp, "propagate")  # costmap_2d's algorithm restated, 1 thread
                line["cpu_baseline"]["ingest_ms"] = 1e3 * (time.perf_counter() - t)
                line["cpu_baseline"]["plan"] = {}
                for name, sc in PLAN_SCENES.items():  # the unmodified reference planner, 1 thread
                    ok, path, it, ms = ref.plan(plan_scene(sc), RES, sc["boundary"], True, sc["start"], sc["goal"], sc["params"])
                    line["cpu_baseline"]["plan"][name] = {"ms": ms, "iterations": it, "path_poses": int(len(path)), "found": bool(ok)}
        except Exception as e:  # the baseline is a report, never a reason to lose the GPU line
            line["cpu_baseline"]["note"] = f"heuristic/teb baseline failed: {e}"

    if rank == 0:
        sys.stdout.flush()
        os.write(result_fd, (json.dumps(line) + "\n").encode())
    out_pinned.free()
    pinned_cost.free()
    h.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
