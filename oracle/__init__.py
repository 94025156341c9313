"""TEST INFRASTRUCTURE ONLY. ctypes access to the two CPU checkers.

* ``ref``  -- oracle/_ref/libpp_ref.so: the reference's own sources compiled unmodified
              (oracle/Makefile); ``None`` where it was never built.
* ``port`` -- oracle/liboracle_port.so: the committed C restatement (oracle/port.c).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline / reference arm may import this
package. The product (path_planner_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


class PlanParams(C.Structure):
    """Search parameters, defaults from src/path_planner/config/params.yaml:32-45."""
    _fields_ = [("motion_type", C.c_int), ("n_theta", C.c_uint32), ("minimum_turning_radius", C.c_float),
                ("cost_penalty", C.c_float), ("change_penalty", C.c_float), ("non_straight_penalty", C.c_float),
                ("reverse_penalty", C.c_float), ("analytic_expansion_ratio", C.c_float),
                ("analytic_expansion_tolerance", C.c_float), ("tolerance", C.c_float), ("max_iterations", C.c_uint32),
                ("max_on_approach_iterations", C.c_uint32)]


def plan_params(motion_type=2, n_theta=72, minimum_turning_radius=0.84, cost_penalty=2.0, change_penalty=0.01,
                non_straight_penalty=1.2, reverse_penalty=3.0, analytic_expansion_ratio=1.0,
                analytic_expansion_tolerance=3.0, tolerance=0.6, max_iterations=100000, max_on_approach_iterations=1000):
    return PlanParams(motion_type, n_theta, minimum_turning_radius, cost_penalty, change_penalty, non_straight_penalty,
                      reverse_penalty, analytic_expansion_ratio, analytic_expansion_tolerance, tolerance, max_iterations,
                      max_on_approach_iterations)


class LayerParams(C.Structure):
    """port_layer_params (oracle/port.h): costmap_2d static + inflation layer configuration."""
    _fields_ = [("track_unknown_space", C.c_int), ("lethal_threshold", C.c_int), ("unknown_cost_value", C.c_int),
                ("trinary_costmap", C.c_int), ("inflate_unknown", C.c_int), ("resolution", C.c_double),
                ("inscribed_radius", C.c_double), ("inflation_radius", C.c_double), ("cost_scaling_factor", C.c_double)]


def layer_params(track_unknown_space=1, lethal_threshold=100, unknown_cost_value=-1, trinary_costmap=1,
                 inflate_unknown=0, resolution=0.05, inscribed_radius=0.25, inflation_radius=0.5,
                 cost_scaling_factor=8.0):
    """Defaults: src/path_planner/config/params.yaml:6,15-22 plus costmap_2d's own defaults."""
    return LayerParams(track_unknown_space, lethal_threshold, unknown_cost_value, trinary_costmap, inflate_unknown,
                       resolution, inscribed_radius, inflation_radius, cost_scaling_factor)


class E2E:
    """oracle/_ref/libpp_e2e.so: the reference search over the drop-in CollisionChecker (see e2e_harness.cpp)."""

    def __init__(self):
        path = os.path.join(_HERE, "_ref", "libpp_e2e.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        self.lib.e2e_heading_fraction.restype = C.c_float
        self.lib.e2e_heading_fraction.argtypes = [C.c_float, C.c_uint32]
        self.lib.e2e_plan.argtypes = [C.c_int, u8p, C.c_uint32, C.c_uint32, C.c_double, f32p, C.c_int, C.c_int, C.c_void_p,
                                      C.POINTER(PlanParams), f32p, f32p, f32p, C.c_int, C.POINTER(C.c_int),
                                      C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                      C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]

        self.lib.e2e_query_log.argtypes = [f32p, u8p, u8p, u8p, C.c_int]

    def heading_fraction(self, start_theta, n_theta):
        return float(self.lib.e2e_heading_fraction(start_theta, n_theta))

    def query_log(self):
        """Footprint questions of the last plan() run with mode + 8: (poses[n,3], kind[n], volume flag[n], volume cost[n])."""
        n = self.lib.e2e_query_log(np.zeros((1, 3), np.float32), np.zeros(1, np.uint8), np.zeros(1, np.uint8), np.zeros(1, np.uint8), 0)
        poses = np.zeros((max(n, 1), 3), np.float32)
        kind, flag, cost = (np.zeros(max(n, 1), np.uint8) for _ in range(3))
        self.lib.e2e_query_log(poses, kind, flag, cost, n)
        return poses[:n], kind[:n], flag[:n], cost[:n]

    def plan(self, mode, cost, res, boundary, hole, start, goal, oracle_volume=None, params=None):
        """mode 0 / 1 (+ 4: analytic-expansion samples at their continuous poses, + 8: log the questions)."""
        P = plan_params(**(params or {}))
        H, W = cost.shape
        b = np.ascontiguousarray(boundary, np.float32)
        path = np.zeros((20000, 3), np.float32)
        n, it = C.c_int(0), C.c_int(0)
        ms, prep = C.c_double(0), C.c_double(0)
        lk, ex = C.c_uint64(0), C.c_uint64(0)
        s = np.ascontiguousarray(start, np.float32)
        g = np.ascontiguousarray(goal, np.float32)
        vol = None if oracle_volume is None else oracle_volume.ctypes.data_as(C.c_void_p)
        ok = self.lib.e2e_plan(mode, cost, W, H, res, b, len(b), int(hole), vol, C.byref(P), s, g, path, len(path),
                               C.byref(n), C.byref(it), C.byref(ms), C.byref(prep), C.byref(lk), C.byref(ex))
        return {"ok": ok, "path": path[:n.value].copy(), "iterations": it.value, "plan_ms": ms.value,
                "prepare_ms": prep.value, "lookups": lk.value, "exact_calls": ex.value}


def load_e2e():
    try:
        return E2E()
    except (FileNotFoundError, OSError):
        return None


def build(verbose=False):
    """Compile the port (always) and _ref (only where /root/reference exists)."""
    r = subprocess.run(["make", "-C", _HERE, "all"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout)


def _load(path):
    return C.CDLL(path) if os.path.exists(path) else None


def _opt(arr, dtype):
    return None if arr is None else arr.ctypes.data_as(C.c_void_p)


class Port:
    def __init__(self):
        path = os.path.join(_HERE, "liboracle_port.so")
        if not os.path.exists(path):
            build()
        self.lib = C.CDLL(path)
        L = self.lib
        L.port_footprint_points.argtypes = [C.c_double, f32p, C.c_int, C.c_int, i32p, C.c_int]
        L.port_se2_volume.argtypes = [u8p, C.c_uint32, C.c_uint32, i32p, C.c_int, C.c_uint32, C.c_float, C.c_float,
                                      C.c_float, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, u8p, C.c_int, C.c_int]
        L.port_footprint_cost_batch.argtypes = [u8p, C.c_uint32, C.c_uint32, i32p, C.c_int, f32p, C.c_uint32, f32p,
                                                u8p, u8p, C.c_int]
        L.port_heuristic.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_float, C.c_float, C.c_float, C.c_float,
                                     C.c_float, C.c_void_p, C.c_uint32, C.c_void_p, f32p]
        L.port_heuristic_fixed_point.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float, f32p]
        L.port_extract_obstacles.argtypes = [u8p, C.c_uint32, C.c_uint32] + [C.c_double] * 8 + [f64p, C.c_uint32]
        L.port_radius_topk.argtypes = [f64p, C.c_uint32, f64p, C.c_uint32, C.c_double, C.c_uint32, u32p, f64p, u32p,
                                       u32p]
        i8p = np.ctypeslib.ndpointer(np.int8, flags="C_CONTIGUOUS")
        lp = C.POINTER(LayerParams)
        L.port_occupancy_from_image.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_double, C.c_double, C.c_int, i8p]
        L.port_static_layer.argtypes = [i8p, C.c_uint32, C.c_uint32, lp, u8p]
        L.port_cell_inflation_radius.argtypes = [lp]
        L.port_cell_inflation_radius.restype = C.c_uint32
        L.port_inflation_cost_table.argtypes = [lp, u8p]
        L.port_inflate_propagate.argtypes = [u8p, C.c_uint32, C.c_uint32, lp]
        L.port_inflate_nearest.argtypes = [u8p, C.c_uint32, C.c_uint32, lp]
        L.port_robot_model.argtypes = [f64p, C.c_int, f64p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.port_obstacle_edges.argtypes = [f64p, C.c_uint32, f64p, u32p, u32p, C.c_uint32, f64p, C.c_int, f64p, f64p,
                                          C.c_void_p]

    def footprint_points(self, resolution, boundary, hole):
        b = np.ascontiguousarray(boundary, np.float32)
        out = np.zeros((4096, 2), np.int32)
        n = self.lib.port_footprint_points(resolution, b, len(b), int(hole), out, 4096)
        assert n <= 4096
        return out[:n].copy()

    def se2_volume(self, cost, pts, n_theta, ox=0.5, oy=0.5, ot=0.0, t0=0, t1=None, y0=0, y1=None, nthreads=0,
                   trig_double=False):
        H, W = cost.shape
        t1 = n_theta if t1 is None else t1
        y1 = H if y1 is None else y1
        pts = np.ascontiguousarray(pts, np.int32)
        out = np.empty((t1 - t0, y1 - y0, W), np.uint8)
        self.lib.port_se2_volume(cost, W, H, pts, len(pts), n_theta, ox, oy, ot, t0, t1, y0, y1, out, nthreads,
                                 int(trig_double))
        return out

    def footprint_cost_batch(self, cost, pts, poses, trig_double=False):
        H, W = cost.shape
        poses = np.ascontiguousarray(poses, np.float32)
        pts = np.ascontiguousarray(pts, np.int32)
        n = len(poses)
        c = np.empty(n, np.float32)
        f = np.empty(n, np.uint8)
        k = np.empty(n, np.uint8)
        self.lib.port_footprint_cost_batch(cost, W, H, pts, len(pts), poses, n, c, f, k, int(trig_double))
        return c, f, k

    def heuristic(self, cost, start, goal, cost_penalty, queries=None):
        H, W = cost.shape
        table = np.empty((H, W), np.float32)
        q = None if queries is None else np.ascontiguousarray(queries, np.float32)
        nq = 0 if q is None else len(q)
        qo = np.empty(nq, np.float32)
        ok = self.lib.port_heuristic(cost, W, H, start[0], start[1], goal[0], goal[1], cost_penalty,
                                     _opt(q, np.float32), nq, qo.ctypes.data_as(C.c_void_p), table)
        return bool(ok), table, qo

    def heuristic_fixed_point(self, cost, goal_xy, cost_penalty):
        H, W = cost.shape
        field = np.empty((H, W), np.float32)
        self.lib.port_heuristic_fixed_point(cost, W, H, int(goal_xy[0]), int(goal_xy[1]), cost_penalty, field)
        return field

    def extract_obstacles(self, cost, res, origin, robot_pose, max_follow_range, behind_ignore_range):
        H, W = cost.shape
        cap = int((cost >= 253).sum()) + 1
        out = np.empty((cap, 2), np.float64)
        n = self.lib.port_extract_obstacles(cost, W, H, res, origin[0], origin[1], robot_pose[0], robot_pose[1],
                                            robot_pose[2], max_follow_range, behind_ignore_range, out, cap)
        return out[:n].copy()

    def radius_topk(self, pts, queries, radius_arg, K):
        pts = np.ascontiguousarray(pts, np.float64)
        queries = np.ascontiguousarray(queries, np.float64)
        Q = len(queries)
        idx = np.zeros((Q, K), np.uint32)
        d2 = np.zeros((Q, K), np.float64)
        cnt = np.zeros(Q, np.uint32)
        found = np.zeros(Q, np.uint32)
        self.lib.port_radius_topk(pts, len(pts), queries, Q, radius_arg, K, idx, d2, cnt, found)
        return idx, d2, cnt, found

    # ---- ingestion (oracle/ingest.c) ----
    def occupancy_from_image(self, image, occ_th, free_th, negate):
        image = np.ascontiguousarray(image, np.uint8)
        H, W = image.shape
        out = np.empty((H, W), np.int8)
        self.lib.port_occupancy_from_image(image, W, H, occ_th, free_th, int(negate), out)
        return out

    def static_layer(self, occ, params):
        occ = np.ascontiguousarray(occ, np.int8)
        H, W = occ.shape
        out = np.empty((H, W), np.uint8)
        self.lib.port_static_layer(occ, W, H, C.byref(params), out)
        return out

    def inflation_cost_table(self, params):
        R = int(self.lib.port_cell_inflation_radius(C.byref(params)))
        tab = np.zeros((R + 2, R + 2), np.uint8)
        self.lib.port_inflation_cost_table(C.byref(params), tab)
        return R, tab

    def inflate(self, static_cost, params, how="nearest"):
        """how = "nearest" (the definition the GPU path is held to) or "propagate" (costmap_2d's queue)."""
        m = np.ascontiguousarray(static_cost, np.uint8).copy()
        H, W = m.shape
        fn = self.lib.port_inflate_nearest if how == "nearest" else self.lib.port_inflate_propagate
        fn(m, W, H, C.byref(params))
        return m

    def costmap_from_occupancy(self, occ, params, how="nearest"):
        return self.inflate(self.static_layer(occ, params), params, how)

    def costmap_window_from_occupancy(self, occ, origin_x, origin_y, win_w, win_h, params, how="nearest"):
        """The follower's rolling window (path_follower/config/params.yaml:12-24): the part of the map under the
        window, unknown beyond the map (costmap_2d leaves the master grid's default there, which is what the static
        layer makes of an unknown cell), inflated from the window's own cells only."""
        occ = np.asarray(occ, np.int8)
        H, W = occ.shape
        win = np.full((win_h, win_w), np.int8(params.unknown_cost_value), np.int8)
        x0, y0 = max(0, origin_x), max(0, origin_y)
        x1, y1 = min(W, origin_x + win_w), min(H, origin_y + win_h)
        if x1 > x0 and y1 > y0:
            win[y0 - origin_y:y1 - origin_y, x0 - origin_x:x1 - origin_x] = occ[y0:y1, x0:x1]
        return self.costmap_from_occupancy(win, params, how)

    @staticmethod
    def window_origin_cell(robot, window_size_m, resolution, map_origin):
        """Costmap2D::updateOrigin (window grid starting at 0, whole cells, truncation toward zero) followed by the
        static layer's worldToMap of the centre of the window's first cell."""
        cell_o = int((robot - window_size_m / 2.0) / resolution)
        grid_o = 0.0 + cell_o * resolution
        return int(np.floor((grid_o + 0.5 * resolution - map_origin) / resolution))

    # ---- obstacle edges (oracle/teb_edges.c) ----
    def robot_model(self, numbers):
        """-> (vertices[nv,2] (empty for a circle), inscribed radius, circumscribed radius)"""
        numbers = np.ascontiguousarray(numbers, np.float64)
        verts = np.zeros((64, 2), np.float64)
        ri, rc = C.c_double(0), C.c_double(0)
        nv = self.lib.port_robot_model(numbers, len(numbers), verts, 64, C.byref(ri), C.byref(rc))
        if nv < 0:
            raise ValueError("bad robot model")
        return verts[:nv].copy(), ri.value, rc.value

    def obstacle_edges(self, poses, pts, idx, cnt, robot, info4, jacobian=True):
        return _obstacle_edges(self.lib.port_obstacle_edges, poses, pts, idx, cnt, robot, info4, jacobian)


def _obstacle_edges(fn, poses, pts, idx, cnt, robot, info4, jacobian):
    poses = np.ascontiguousarray(poses, np.float64)
    pts = np.ascontiguousarray(pts, np.float64)
    idx = np.ascontiguousarray(idx, np.uint32)
    cnt = np.ascontiguousarray(cnt, np.uint32)
    robot = np.ascontiguousarray(robot, np.float64)
    info4 = np.ascontiguousarray(info4, np.float64)
    Q, K = idx.shape
    err = np.zeros((Q, K, 2), np.float64)
    jac = np.zeros((Q, K, 2, 3), np.float64) if jacobian else None
    rc = fn(poses, Q, pts, idx, cnt, K, robot, len(robot), info4, err, _opt(jac, np.float64))
    if rc != 0:
        raise ValueError("bad robot model")
    return err, jac


class TebRef:
    """oracle/_ref/libpp_teb_ref.so: the reference's own robot model and EdgeObstacle (teb_ref_harness.cpp)."""

    def __init__(self):
        path = os.path.join(_HERE, "_ref", "libpp_teb_ref.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        self.lib.teb_ref_robot_radii.argtypes = [f64p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        self.lib.teb_ref_obstacle_edges.argtypes = [f64p, C.c_uint32, f64p, u32p, u32p, C.c_uint32, f64p, C.c_int, f64p,
                                                    f64p, C.c_void_p]

    def robot_radii(self, numbers):
        numbers = np.ascontiguousarray(numbers, np.float64)
        ri, rc = C.c_double(0), C.c_double(0)
        self.lib.teb_ref_robot_radii(numbers, len(numbers), C.byref(ri), C.byref(rc))
        return ri.value, rc.value

    def obstacle_edges(self, poses, pts, idx, cnt, robot, info4, jacobian=True):
        return _obstacle_edges(self.lib.teb_ref_obstacle_edges, poses, pts, idx, cnt, robot, info4, jacobian)


def load_teb_ref():
    try:
        return TebRef()
    except (FileNotFoundError, OSError):
        return None


class Ref:
    """The reference's own code. Footprints are given as the reference takes them: the
    boundary vector (1, 2 or 4 numbers, metres), the costmap resolution and the hole flag."""

    def __init__(self, double_trig=False):
        name = "libpp_ref_dtrig.so" if double_trig else "libpp_ref.so"
        path = os.path.join(_HERE, "_ref", name)
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        L = self.lib
        L.ref_footprint_points.argtypes = [C.c_double, f32p, C.c_int, C.c_int, i32p, C.c_int]
        L.ref_se2_volume.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_double, f32p, C.c_int, C.c_int, C.c_uint32,
                                     C.c_float, C.c_float, C.c_float, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                     u8p, C.c_int]
        L.ref_footprint_cost_batch.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_double, f32p, C.c_int, C.c_int, f32p,
                                               C.c_uint32, f32p, u8p, u8p]
        L.ref_heuristic.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_double, C.c_float, C.c_float, C.c_float,
                                    C.c_float, C.c_float, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, f32p,
                                    C.POINTER(C.c_double)]
        L.ref_radius_topk.argtypes = [f64p, C.c_uint32, f64p, C.c_uint32, C.c_double, C.c_uint32, u32p, f64p, u32p,
                                      u32p, C.POINTER(C.c_double), C.POINTER(C.c_double)]

    def footprint_points(self, resolution, boundary, hole):
        b = np.ascontiguousarray(boundary, np.float32)
        out = np.zeros((4096, 2), np.int32)
        n = self.lib.ref_footprint_points(resolution, b, len(b), int(hole), out, 4096)
        assert n <= 4096
        return out[:n].copy()

    def se2_volume(self, cost, res, boundary, hole, n_theta, ox=0.5, oy=0.5, ot=0.0, t0=0, t1=None, y0=0, y1=None,
                   nthreads=0):
        H, W = cost.shape
        t1 = n_theta if t1 is None else t1
        y1 = H if y1 is None else y1
        b = np.ascontiguousarray(boundary, np.float32)
        out = np.empty((t1 - t0, y1 - y0, W), np.uint8)
        self.lib.ref_se2_volume(cost, W, H, res, b, len(b), int(hole), n_theta, ox, oy, ot, t0, t1, y0, y1, out,
                                nthreads)
        return out

    def footprint_cost_batch(self, cost, res, boundary, hole, poses):
        H, W = cost.shape
        b = np.ascontiguousarray(boundary, np.float32)
        poses = np.ascontiguousarray(poses, np.float32)
        n = len(poses)
        c = np.empty(n, np.float32)
        f = np.empty(n, np.uint8)
        k = np.empty(n, np.uint8)
        self.lib.ref_footprint_cost_batch(cost, W, H, res, b, len(b), int(hole), poses, n, c, f, k)
        return c, f, k

    def heuristic(self, cost, res, start, goal, cost_penalty, queries=None, n_theta=72):
        H, W = cost.shape
        table = np.empty((H, W), np.float32)
        q = None if queries is None else np.ascontiguousarray(queries, np.float32)
        nq = 0 if q is None else len(q)
        qo = np.empty(nq, np.float32)
        sec = C.c_double(0)
        ok = self.lib.ref_heuristic(cost, W, H, res, start[0], start[1], goal[0], goal[1], cost_penalty, n_theta,
                                    _opt(q, np.float32), nq, qo.ctypes.data_as(C.c_void_p), table, C.byref(sec))
        self.last_seconds = sec.value
        return bool(ok), table, qo

    def radius_topk(self, pts, queries, radius_arg, K):
        pts = np.ascontiguousarray(pts, np.float64)
        queries = np.ascontiguousarray(queries, np.float64)
        Q = len(queries)
        idx = np.zeros((Q, K), np.uint32)
        d2 = np.zeros((Q, K), np.float64)
        cnt = np.zeros(Q, np.uint32)
        found = np.zeros(Q, np.uint32)
        bs, qs = C.c_double(0), C.c_double(0)
        self.lib.ref_radius_topk(pts, len(pts), queries, Q, radius_arg, K, idx, d2, cnt, found, C.byref(bs),
                                 C.byref(qs))
        self.last_build_seconds, self.last_query_seconds = bs.value, qs.value
        return idx, d2, cnt, found

    def plan(self, cost, res, boundary, hole, start, goal, params=None):
        """Unmodified reference planner (AstarHybrid::createPath). Returns (ok, path[n,3], iterations, ms)."""
        P = plan_params(**(params or {}))
        H, W = cost.shape
        b = np.ascontiguousarray(boundary, np.float32)
        path = np.zeros((20000, 3), np.float32)
        n, it, ms = C.c_int(0), C.c_int(0), C.c_double(0)
        s = np.ascontiguousarray(start, np.float32)
        g = np.ascontiguousarray(goal, np.float32)
        self.lib.ref_plan.argtypes = [u8p, C.c_uint32, C.c_uint32, C.c_double, f32p, C.c_int, C.c_int, C.POINTER(PlanParams),
                                      f32p, f32p, f32p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double)]
        ok = self.lib.ref_plan(cost, W, H, res, b, len(b), int(hole), C.byref(P), s, g, path, len(path), C.byref(n),
                               C.byref(it), C.byref(ms))
        return bool(ok), path[:n.value].copy(), it.value, ms.value


def load_port():
    return Port()


def load_ref(double_trig=False):
    """Returns None where oracle/_ref was never built (no /root/reference on that machine)."""
    try:
        return Ref(double_trig)
    except FileNotFoundError:
        return None
