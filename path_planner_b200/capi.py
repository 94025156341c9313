"""ctypes binding of libpp_b200.so (include/pp_b200.h) for the test-suite and bench.py.

This is harness code: the product is the shared library and the C++ host classes in
path_planner_b200/host/. There is no fallback of any kind -- if the library is missing, or no
CUDA device is present, loading or pp_create fails loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PP_B200_LIB", os.path.join(_HERE, "libpp_b200.so"))  # override: A/B runs of two builds

# every symbol include/pp_b200.h declares
SYMBOLS = [
    "pp_version", "pp_create", "pp_destroy", "pp_last_error", "pp_sync", "pp_host_alloc", "pp_host_free",
    "pp_timer_start", "pp_timer_stop", "pp_l2_flush", "pp_launch_count", "pp_costmap_upload", "pp_costmap_update_rect", "pp_costmap_download",
    "pp_footprint_from_boundary", "pp_footprint_set", "pp_footprint_get", "pp_se2_volume_build", "pp_se2_volume_build_bins", "pp_se2_volume_build_sharded",
    "pp_se2_volume_refresh", "pp_se2_volume_download", "pp_se2_volume_download_rows", "pp_se2_volume_ptr", "pp_se2_volume_set_mode",
    "pp_se2_volume_stats", "pp_footprint_cost_batch", "pp_heuristic_build", "pp_heuristic_ptr", "pp_heuristic_stats", "pp_heuristic_build_noalloc", "pp_plan", "pp_plan_stats", "pp_plan_set_query_log", "pp_plan_query_log",
    "pp_obstacles_extract", "pp_radius_topk", "pp_comm_unique_id", "pp_comm_init", "pp_se2_volume_allgather", "pp_se2_volume_reserve",
    "pp_se2_volume_ipc_export", "pp_se2_volume_ipc_open", "pp_se2_volume_set_push", "pp_se2_volume_set_rows",
    "pp_occupancy_from_image", "pp_occupancy_upload", "pp_occupancy_update_rect", "pp_occupancy_download",
    "pp_costmap_from_occupancy", "pp_robot_model_set", "pp_obstacle_edges",
    "pp_costmap_window_from_occupancy", "pp_window_origin_cell",
]

_vp = C.c_void_p
_u32 = C.c_uint32
_lib = None


class PlanParams(C.Structure):
    """pp_plan_params; defaults from the reference's config/params.yaml:32-45."""
    _fields_ = [("motion_type", C.c_int32), ("n_theta", C.c_uint32), ("minimum_turning_radius", C.c_float),
                ("cost_penalty", C.c_float), ("change_penalty", C.c_float), ("non_straight_penalty", C.c_float),
                ("reverse_penalty", C.c_float), ("analytic_expansion_ratio", C.c_float),
                ("analytic_expansion_tolerance", C.c_float), ("tolerance", C.c_float), ("max_iterations", C.c_uint32),
                ("max_on_approach_iterations", C.c_uint32), ("flags", C.c_uint32)]

    @classmethod
    def make(cls, motion_type=2, n_theta=72, minimum_turning_radius=0.84, cost_penalty=2.0, change_penalty=0.01,
             non_straight_penalty=1.2, reverse_penalty=3.0, analytic_expansion_ratio=1.0,
             analytic_expansion_tolerance=3.0, tolerance=0.6, max_iterations=100000, max_on_approach_iterations=1000,
             exact_analytic=False, validate_path=False):
        return cls(motion_type, n_theta, minimum_turning_radius, cost_penalty, change_penalty, non_straight_penalty,
                   reverse_penalty, analytic_expansion_ratio, analytic_expansion_tolerance, tolerance, max_iterations,
                   max_on_approach_iterations, (1 if exact_analytic else 0) | (2 if validate_path else 0))


class PlanStats(C.Structure):
    """pp_plan_stats_t"""
    _fields_ = [("lookups", C.c_uint64), ("analytic_attempts", C.c_uint32), ("analytic_confirm_batches", C.c_uint32),
                ("exact_host_checks", C.c_uint32), ("false_collisions_overruled", C.c_uint32), ("false_free_caught", C.c_uint32),
                ("path_poses_checked", C.c_uint32), ("path_poses_in_collision", C.c_uint32), ("replans", C.c_uint32)]


class LayerParams(C.Structure):
    """pp_layer_params; defaults from the reference's config/params.yaml:6,15-22 and costmap_2d's own."""
    _fields_ = [("track_unknown_space", C.c_int), ("lethal_threshold", C.c_int), ("unknown_cost_value", C.c_int),
                ("trinary_costmap", C.c_int), ("inflate_unknown", C.c_int), ("resolution", C.c_double),
                ("inscribed_radius", C.c_double), ("inflation_radius", C.c_double), ("cost_scaling_factor", C.c_double)]

    @classmethod
    def make(cls, track_unknown_space=1, lethal_threshold=100, unknown_cost_value=-1, trinary_costmap=1,
             inflate_unknown=0, resolution=0.05, inscribed_radius=0.25, inflation_radius=0.5, cost_scaling_factor=8.0):
        return cls(track_unknown_space, lethal_threshold, unknown_cost_value, trinary_costmap, inflate_unknown,
                   resolution, inscribed_radius, inflation_radius, cost_scaling_factor)


class ObstacleEdgeParams(C.Structure):
    """pp_obstacle_edge_params; defaults from FollowerInfo (teb_path_follower/commons.h:219-222)."""
    _fields_ = [("obstacle_collision_dist", C.c_double), ("margin_obstacle_collision", C.c_double),
                ("obstacle_inflation_dist", C.c_double), ("margin_obstacle_inflation", C.c_double)]

    @classmethod
    def make(cls, obstacle_collision_dist=0.1, margin_obstacle_collision=0.05, obstacle_inflation_dist=0.2,
             margin_obstacle_inflation=0.1):
        return cls(obstacle_collision_dist, margin_obstacle_collision, obstacle_inflation_dist, margin_obstacle_inflation)


class PPError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"pp_b200 status {status}: {message}")
        self.status = status


def load():
    """Load libpp_b200.so. Raises if it has not been built (python -c 'import __graft_entry__ as g; g.build()')."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it first (make -C path_planner_b200/csrc); "
                          "there is no CPU or PyTorch fallback for this library")
    L = C.CDLL(LIB_PATH)
    L.pp_version.restype = C.c_char_p
    L.pp_last_error.restype = C.c_char_p
    L.pp_last_error.argtypes = [_vp]
    L.pp_create.argtypes = [C.POINTER(_vp), C.c_int]
    L.pp_destroy.argtypes = [_vp]
    L.pp_sync.argtypes = [_vp]
    L.pp_host_alloc.argtypes = [C.POINTER(_vp), C.c_size_t]
    L.pp_host_free.argtypes = [_vp]
    L.pp_timer_start.argtypes = [_vp]
    L.pp_timer_stop.argtypes = [_vp, C.POINTER(C.c_float)]
    L.pp_l2_flush.argtypes = [_vp, C.c_size_t]
    L.pp_launch_count.argtypes = [_vp, C.POINTER(C.c_uint64)]
    L.pp_costmap_upload.argtypes = [_vp, _vp, _u32, _u32, C.c_int]
    L.pp_costmap_update_rect.argtypes = [_vp, _u32, _u32, _u32, _u32, _vp, C.c_size_t]
    L.pp_costmap_download.argtypes = [_vp, _vp]
    L.pp_footprint_from_boundary.argtypes = [_vp, C.c_double, _vp, C.c_int, C.c_int]
    L.pp_footprint_set.argtypes = [_vp, _vp, _u32]
    L.pp_footprint_get.argtypes = [_vp, _vp, _u32, C.POINTER(_u32)]
    L.pp_se2_volume_build.argtypes = [_vp, _u32, C.c_float, C.c_float, C.c_float, _u32, _u32]
    L.pp_se2_volume_build_bins.argtypes = [_vp, _u32, C.c_float, C.c_float, C.c_float, _vp, _u32]
    L.pp_se2_volume_build_sharded.argtypes = [_vp, _u32, C.c_float, C.c_float, C.c_float, _u32, _u32]
    L.pp_se2_volume_refresh.argtypes = [_vp]
    L.pp_se2_volume_download.argtypes = [_vp, _u32, _u32, _vp, C.c_int]
    L.pp_se2_volume_download_rows.argtypes = [_vp, _u32, _u32, _u32, _u32, _vp]
    L.pp_se2_volume_ptr.argtypes = [_vp, C.POINTER(_vp), C.POINTER(C.c_uint64)]
    L.pp_se2_volume_set_mode.argtypes = [_vp, C.c_int]
    L.pp_se2_volume_stats.argtypes = [_vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_double)]
    L.pp_footprint_cost_batch.argtypes = [_vp, _vp, _u32, _vp, _vp]
    L.pp_heuristic_build.argtypes = [_vp, _u32, _u32, C.c_float, _vp]
    L.pp_heuristic_build_noalloc.argtypes = [_vp, _u32, _u32, _u32, _u32, C.c_float, _vp, _vp]
    L.pp_plan.argtypes = [_vp, C.POINTER(PlanParams), C.c_double, C.c_double, C.c_double, _vp, _vp, _vp, _u32,
                          C.POINTER(_u32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_double),
                          C.POINTER(C.c_double)]
    L.pp_plan_stats.argtypes = [_vp, C.POINTER(PlanStats)]
    L.pp_plan_set_query_log.argtypes = [_vp, C.c_int]
    L.pp_plan_query_log.argtypes = [_vp, _vp, _vp, _vp, _vp, _u32, C.POINTER(_u32)]
    L.pp_heuristic_ptr.argtypes = [_vp, C.POINTER(_vp)]
    L.pp_heuristic_stats.argtypes = [_vp, C.POINTER(_u32), C.POINTER(C.c_uint64)]
    L.pp_obstacles_extract.argtypes = [_vp] + [C.c_double] * 8 + [_vp, _u32, C.POINTER(_u32)]
    L.pp_radius_topk.argtypes = [_vp, _vp, _u32, _vp, _u32, C.c_double, _u32, _vp, _vp, _vp]
    L.pp_comm_unique_id.argtypes = [_vp]
    L.pp_comm_init.argtypes = [_vp, _vp, C.c_int, C.c_int]
    L.pp_se2_volume_allgather.argtypes = [_vp]
    L.pp_se2_volume_reserve.argtypes = [_vp, _u32]
    L.pp_se2_volume_ipc_export.argtypes = [_vp, _vp]
    L.pp_se2_volume_ipc_open.argtypes = [_vp, _vp, C.c_int, C.c_int]
    L.pp_se2_volume_set_push.argtypes = [_vp, C.c_int]
    L.pp_se2_volume_set_rows.argtypes = [_vp, _u32, _u32]
    L.pp_occupancy_from_image.argtypes = [_vp, _vp, _u32, _u32, C.c_double, C.c_double, C.c_int]
    L.pp_occupancy_upload.argtypes = [_vp, _vp, _u32, _u32]
    L.pp_occupancy_update_rect.argtypes = [_vp, _u32, _u32, _u32, _u32, _vp, C.c_size_t]
    L.pp_occupancy_download.argtypes = [_vp, _vp]
    L.pp_costmap_from_occupancy.argtypes = [_vp, C.POINTER(LayerParams)]
    L.pp_costmap_window_from_occupancy.argtypes = [_vp, _vp, C.c_int32, C.c_int32, _u32, _u32, C.POINTER(LayerParams)]
    L.pp_window_origin_cell.argtypes = [C.c_double, C.c_double, C.c_double, C.c_double]
    L.pp_window_origin_cell.restype = C.c_int32
    L.pp_robot_model_set.argtypes = [_vp, _vp, _u32, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.pp_obstacle_edges.argtypes = [_vp, _vp, _u32, _u32, C.POINTER(ObstacleEdgeParams), _vp, _vp]
    _lib = L
    return L


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_vp)


class PinnedBuffer:
    """numpy view of page-locked memory from pp_host_alloc."""

    def __init__(self, nbytes):
        L = load()
        p = _vp()
        st = L.pp_host_alloc(C.byref(p), nbytes)
        if st != 0:
            raise PPError(st, "pp_host_alloc failed")
        self._p = p
        self.nbytes = nbytes
        self.array = np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(p.value))

    def free(self):
        if self._p:
            load().pp_host_free(self._p)
            self._p = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Handle:
    """One pp_handle. Method names follow the C entry points without the pp_ prefix."""

    def __init__(self, device=0):
        self.L = load()
        self.h = _vp()
        st = self.L.pp_create(C.byref(self.h), device)
        if st != 0:
            raise PPError(st, "pp_create failed (no CUDA device? this library has no CPU path)")
        self.W = self.H = 0
        self.n_theta = 0

    def close(self):
        if self.h:
            self.L.pp_destroy(self.h)
            self.h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, st):
        if st != 0:
            raise PPError(st, self.L.pp_last_error(self.h).decode())

    def sync(self):
        self._ck(self.L.pp_sync(self.h))

    def timer_start(self):
        self._ck(self.L.pp_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float(0)
        self._ck(self.L.pp_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def l2_flush(self, nbytes=256 << 20):
        self._ck(self.L.pp_l2_flush(self.h, nbytes))

    def launch_count(self):
        n = C.c_uint64(0)
        self._ck(self.L.pp_launch_count(self.h, C.byref(n)))
        return n.value

    # K0
    def costmap_upload(self, cost, pinned=False):
        assert cost.dtype == np.uint8 and cost.ndim == 2 and cost.flags.c_contiguous
        self.H, self.W = cost.shape
        self._ck(self.L.pp_costmap_upload(self.h, _ptr(cost), self.W, self.H, int(pinned)))

    def costmap_update_rect(self, x0, y0, patch):
        assert patch.dtype == np.uint8 and patch.ndim == 2
        patch = np.ascontiguousarray(patch)
        hgt, w = patch.shape
        self._ck(self.L.pp_costmap_update_rect(self.h, x0, y0, w, hgt, _ptr(patch), w))

    def costmap_download(self):
        out = np.empty((self.H, self.W), np.uint8)
        self._ck(self.L.pp_costmap_download(self.h, _ptr(out)))
        return out

    # N2 ingestion
    def occupancy_from_image(self, image, occupied_thresh, free_thresh, negate):
        assert image.dtype == np.uint8 and image.ndim == 2 and image.flags.c_contiguous
        self.occ_H, self.occ_W = image.shape
        self._ck(self.L.pp_occupancy_from_image(self.h, _ptr(image), self.occ_W, self.occ_H, occupied_thresh,
                                                free_thresh, int(negate)))

    def occupancy_upload(self, occ):
        assert occ.dtype == np.int8 and occ.ndim == 2 and occ.flags.c_contiguous
        self.occ_H, self.occ_W = occ.shape
        self._ck(self.L.pp_occupancy_upload(self.h, _ptr(occ), self.occ_W, self.occ_H))

    def occupancy_update_rect(self, x0, y0, patch):
        assert patch.dtype == np.int8 and patch.ndim == 2
        patch = np.ascontiguousarray(patch)
        hgt, w = patch.shape
        self._ck(self.L.pp_occupancy_update_rect(self.h, x0, y0, w, hgt, _ptr(patch), w))

    def occupancy_download(self):
        out = np.empty((self.occ_H, self.occ_W), np.int8)
        self._ck(self.L.pp_occupancy_download(self.h, _ptr(out)))
        return out

    def costmap_from_occupancy(self, params=None):
        p = params if params is not None else LayerParams.make()
        self._ck(self.L.pp_costmap_from_occupancy(self.h, C.byref(p)))
        self.H, self.W = self.occ_H, self.occ_W

    def costmap_window_from_occupancy(self, src, origin_x, origin_y, win_w, win_h, params=None):
        """Rolling window: this handle's costmap becomes the win_w x win_h window over `src`'s occupancy grid."""
        p = params if params is not None else LayerParams.make()
        self._ck(self.L.pp_costmap_window_from_occupancy(self.h, src.h, int(origin_x), int(origin_y), int(win_w), int(win_h), C.byref(p)))
        self.H, self.W = int(win_h), int(win_w)

    # footprint
    def footprint_from_boundary(self, resolution, boundary, hole):
        b = np.ascontiguousarray(boundary, np.float32)
        self._ck(self.L.pp_footprint_from_boundary(self.h, resolution, _ptr(b), len(b), int(hole)))

    def footprint_set(self, offsets):
        o = np.ascontiguousarray(offsets, np.int32)
        self._ck(self.L.pp_footprint_set(self.h, _ptr(o), len(o)))

    def footprint_get(self):
        n = _u32(0)
        self._ck(self.L.pp_footprint_get(self.h, None, 0, C.byref(n)))
        out = np.zeros((n.value, 2), np.int32)
        self._ck(self.L.pp_footprint_get(self.h, _ptr(out), n.value, C.byref(n)))
        return out

    # K1
    def se2_volume_build_bins(self, n_theta, bins, ox=0.5, oy=0.5, ot=0.0):
        b = np.ascontiguousarray(bins, np.uint32)
        self._ck(self.L.pp_se2_volume_build_bins(self.h, n_theta, ox, oy, ot, b.ctypes.data, len(b)))
        self.n_theta = n_theta

    def se2_volume_build_sharded(self, n_theta, n_chunks=3, push=False, exchange_only=False, ox=0.5, oy=0.5, ot=0.0):
        self._ck(self.L.pp_se2_volume_build_sharded(self.h, n_theta, ox, oy, ot, n_chunks, (1 if push else 0) | (2 if exchange_only else 0)))
        self.n_theta = n_theta

    def se2_volume_build(self, n_theta, ox=0.5, oy=0.5, ot=0.0, t_begin=0, t_end=None):
        t_end = n_theta if t_end is None else t_end
        self.n_theta = n_theta
        self._ck(self.L.pp_se2_volume_build(self.h, n_theta, ox, oy, ot, t_begin, t_end))

    def se2_volume_refresh(self):
        self._ck(self.L.pp_se2_volume_refresh(self.h))

    def se2_volume_download(self, t_begin=0, t_end=None, out=None, pinned=False):
        t_end = self.n_theta if t_end is None else t_end
        if out is None:
            out = np.empty((t_end - t_begin, self.H, self.W), np.uint8)
        self._ck(self.L.pp_se2_volume_download(self.h, t_begin, t_end, _ptr(out), int(pinned)))
        return out

    def se2_volume_download_rows(self, y_begin, y_end, t_begin=0, t_end=None, out=None):
        t_end = self.n_theta if t_end is None else t_end
        if out is None:
            out = np.empty((t_end - t_begin, y_end - y_begin, self.W), np.uint8)
        self._ck(self.L.pp_se2_volume_download_rows(self.h, t_begin, t_end, y_begin, y_end, _ptr(out)))
        return out

    def se2_volume_ptr(self):
        p = _vp()
        n = C.c_uint64(0)
        self._ck(self.L.pp_se2_volume_ptr(self.h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def se2_volume_set_mode(self, mode):
        self._ck(self.L.pp_se2_volume_set_mode(self.h, mode))

    def se2_volume_stats(self):
        a, b, s = C.c_uint64(0), C.c_uint64(0), C.c_double(0)
        self._ck(self.L.pp_se2_volume_stats(self.h, C.byref(a), C.byref(b), C.byref(s)))
        return {"fast_tiles": a.value, "generic_tiles": b.value, "samples_per_cell": s.value}

    # K1b
    def footprint_cost_batch(self, poses):
        poses = np.ascontiguousarray(poses, np.float32)
        n = len(poses)
        cost = np.zeros(n, np.float32)
        flag = np.zeros(n, np.uint8)
        self._ck(self.L.pp_footprint_cost_batch(self.h, _ptr(poses), n, _ptr(cost), _ptr(flag)))
        return cost, flag

    # K2
    def heuristic_build(self, goal_x, goal_y, cost_penalty, download=True):
        out = np.empty((self.H, self.W), np.float32) if download else None
        self._ck(self.L.pp_heuristic_build(self.h, goal_x, goal_y, cost_penalty, _ptr(out)))
        return out

    def heuristic_stats(self):
        r, v = _u32(0), C.c_uint64(0)
        self._ck(self.L.pp_heuristic_stats(self.h, C.byref(r), C.byref(v)))
        return {"rounds": r.value, "tile_visits": v.value}

    # N1
    def plan(self, resolution, origin, start, goal, params=None, max_points=20000):
        P = PlanParams.make(**(params or {}))
        s = np.ascontiguousarray(start, np.float32)
        g = np.ascontiguousarray(goal, np.float32)
        path = np.zeros((max_points, 3), np.float32)
        n, it, found = _u32(0), C.c_int32(0), C.c_int32(0)
        t_ms, s_ms = C.c_double(0), C.c_double(0)
        self._ck(self.L.pp_plan(self.h, C.byref(P), resolution, origin[0], origin[1], _ptr(s), _ptr(g), _ptr(path), max_points,
                                C.byref(n), C.byref(it), C.byref(found), C.byref(t_ms), C.byref(s_ms)))
        st = PlanStats()
        self._ck(self.L.pp_plan_stats(self.h, C.byref(st)))
        return {"ok": found.value, "path": path[:n.value].copy(), "iterations": it.value, "tables_ms": t_ms.value,
                "search_ms": s_ms.value, "stats": {f: int(getattr(st, f)) for f, _ in PlanStats._fields_}}

    def plan_set_query_log(self, enable):
        self._ck(self.L.pp_plan_set_query_log(self.h, int(enable)))

    def plan_query_log(self):
        """(poses[n,3] float32 (cells, cells, rad), kind[n], volume flag[n], volume cost[n]) of the last pp_plan."""
        n = _u32(0)
        self._ck(self.L.pp_plan_query_log(self.h, None, None, None, None, 0, C.byref(n)))
        poses = np.zeros((n.value, 3), np.float32)
        kind, flag, cost = (np.zeros(n.value, np.uint8) for _ in range(3))
        if n.value:
            self._ck(self.L.pp_plan_query_log(self.h, _ptr(poses), _ptr(kind), _ptr(flag), _ptr(cost), n.value, C.byref(n)))
        return poses, kind, flag, cost

    # K3
    def obstacles_extract(self, resolution, origin, robot_pose, max_follow_range, behind_ignore_range, download=True):
        n = _u32(0)
        self._ck(self.L.pp_obstacles_extract(self.h, resolution, origin[0], origin[1], robot_pose[0], robot_pose[1],
                                             robot_pose[2], max_follow_range, behind_ignore_range, None, 0, C.byref(n)))
        if not download:
            return n.value
        out = np.empty((n.value, 2), np.float64)
        if n.value:
            self._ck(self.L.pp_obstacles_extract(self.h, resolution, origin[0], origin[1], robot_pose[0], robot_pose[1],
                                                 robot_pose[2], max_follow_range, behind_ignore_range, _ptr(out),
                                                 n.value, C.byref(n)))
        return out

    def radius_topk(self, pts, queries, radius_arg, K):
        queries = np.ascontiguousarray(queries, np.float64)
        Q = len(queries)
        N = 0
        if pts is not None:
            pts = np.ascontiguousarray(pts, np.float64)
            N = len(pts)
        idx = np.zeros((Q, K), np.uint32)
        d2 = np.zeros((Q, K), np.float64)
        cnt = np.zeros(Q, np.uint32)
        self._ck(self.L.pp_radius_topk(self.h, _ptr(pts), N, _ptr(queries), Q, radius_arg, K, _ptr(idx), _ptr(d2),
                                       _ptr(cnt)))
        return idx, d2, cnt

    # N3
    def robot_model_set(self, numbers):
        """-> (inscribed radius, circumscribed radius)"""
        numbers = np.ascontiguousarray(numbers, np.float64)
        ri, rc = C.c_double(0), C.c_double(0)
        self._ck(self.L.pp_robot_model_set(self.h, _ptr(numbers), len(numbers), C.byref(ri), C.byref(rc)))
        return ri.value, rc.value

    def obstacle_edges(self, poses_xyt, K, params=None, jacobian=True):
        """Errors [Q,K,2] and Jacobians [Q,K,2,3] for the association of the last radius_topk."""
        poses = np.ascontiguousarray(poses_xyt, np.float64)
        Q = len(poses)
        p = params if params is not None else ObstacleEdgeParams.make()
        err = np.zeros((Q, K, 2), np.float64)
        jac = np.zeros((Q, K, 2, 3), np.float64) if jacobian else None
        self._ck(self.L.pp_obstacle_edges(self.h, _ptr(poses), Q, K, C.byref(p), _ptr(err), _ptr(jac)))
        return err, jac

    # multi-GPU
    def comm_init(self, uid_bytes, rank, world):
        buf = C.create_string_buffer(bytes(uid_bytes), 128)
        self._ck(self.L.pp_comm_init(self.h, buf, rank, world))

    def se2_volume_allgather(self):
        self._ck(self.L.pp_se2_volume_allgather(self.h))

    def se2_volume_reserve(self, n_theta):
        self.n_theta = n_theta
        self._ck(self.L.pp_se2_volume_reserve(self.h, n_theta))

    def se2_volume_ipc_export(self):
        buf = C.create_string_buffer(64)
        self._ck(self.L.pp_se2_volume_ipc_export(self.h, buf))
        return buf.raw

    def se2_volume_ipc_open(self, handles, rank, world):
        blob = b"".join(handles)
        assert len(blob) == 64 * world
        self._ck(self.L.pp_se2_volume_ipc_open(self.h, C.create_string_buffer(blob, len(blob)), rank, world))

    def se2_volume_set_push(self, enable):
        self._ck(self.L.pp_se2_volume_set_push(self.h, int(enable)))

    def se2_volume_set_rows(self, y_begin, y_end):
        self._ck(self.L.pp_se2_volume_set_rows(self.h, y_begin, y_end))


def heuristic_build_noalloc(d_cost_ptr, W, H, goal_x, goal_y, cost_penalty, d_field_ptr, stream=None):
    """pp_heuristic_build_noalloc: device pointers in and out (integers), no handle, no allocation."""
    st = load().pp_heuristic_build_noalloc(d_cost_ptr, W, H, goal_x, goal_y, cost_penalty, d_field_ptr, stream)
    if st != 0:
        raise PPError(st, "pp_heuristic_build_noalloc failed (message on stderr)")


def comm_unique_id():
    buf = C.create_string_buffer(128)
    st = load().pp_comm_unique_id(buf)
    if st != 0:
        raise PPError(st, "pp_comm_unique_id failed")
    return buf.raw
