"""dev probe: host search time on the bench's plan scenes, tables from the CPU oracle (no GPU needed).
    python tools/search_probe.py [save|check]   -- save / compare paths in /tmp/probe/paths.npz"""
import ctypes as C, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle
import bench
so = "/tmp/probe/libsearch_probe.so"
subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-O3", "-DNDEBUG", "-fPIC", "-ffp-contract=off", "-shared", "-o", so,
                       os.path.join(ROOT, "tools/search_probe.cpp"), os.path.join(ROOT, "path_planner_b200/host/hybrid_astar.cpp")])
L = C.CDLL(so)
L.probe_heading_fraction.restype = C.c_float
L.probe_heading_fraction.argtypes = [C.c_int, C.c_uint, C.c_uint, C.c_uint, C.c_double, C.c_float]
port = oracle.load_port()
mode = sys.argv[1] if len(sys.argv) > 1 else ""
saved = dict(np.load("/tmp/probe/paths.npz")) if mode == "check" else {}
out = {}
for name, sc in bench.PLAN_SCENES.items():
    cache = f"/tmp/probe/{name}.npz"
    cost = bench.plan_scene(sc)
    Wd = sc["W"]
    mt = sc["params"].get("motion_type", 2)
    frac = L.probe_heading_fraction(mt, 72, Wd, Wd, bench.RES, sc["start"][2])
    if os.path.exists(cache):
        z = np.load(cache); vol, field = z["vol"], z["field"]
    else:
        pts = port.footprint_points(bench.RES, sc["boundary"], True)
        vol = port.se2_volume(cost, pts, 72, 0.5, 0.5, frac)
        gx, gy = int(sc["goal"][0] / bench.RES), int(sc["goal"][1] / bench.RES)
        field = port.heuristic_fixed_point(cost, (gx, gy), 2.0)
        np.savez(cache, vol=vol, field=field)
    path = np.zeros((20000, 3), np.float32)
    n, it, ms = C.c_int(0), C.c_int(0), C.c_double(0)
    s = np.array(sc["start"], np.float32); g = np.array(sc["goal"], np.float32)
    ok = L.probe_plan(vol.ctypes.data_as(C.c_void_p), field.ctypes.data_as(C.c_void_p), Wd, Wd, 72, C.c_float(frac), C.c_double(bench.RES),
                      s.ctypes.data_as(C.c_void_p), g.ctypes.data_as(C.c_void_p), mt, sc["params"].get("max_iterations", 100000),
                      path.ctypes.data_as(C.c_void_p), 20000, C.byref(n), C.byref(it), C.byref(ms), 3)
    print(f"{name}: ok={ok} iterations={it.value} poses={n.value} search {ms.value:.1f} ms")
    out[name] = path[:n.value].copy(); out[name + "_it"] = np.array(it.value)
    if mode == "check":
        assert np.array_equal(saved[name], out[name]) and saved[name + "_it"] == it.value, "PATH CHANGED"
if mode == "save":
    np.savez("/tmp/probe/paths.npz", **out)
if mode == "check":
    print("paths and iteration counts unchanged")
