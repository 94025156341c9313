"""N4: the nav_core::BaseGlobalPlanner adapter (path_planner_b200/host/ros/), compiled against the stand-in
ROS headers of tests/ros_shim (ROS is not installed here). CPU: it compiles, links against the C ABI and
exports the harness entry point. GPU: makePlan through the plugin interface returns the plan pp_plan returns."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, RES

SO = os.path.join(ROOT, "tests", "ros_shim", "libpp_ros_adapter_test.so")


def build_adapter():
    lib_dir = os.path.join(ROOT, "path_planner_b200")
    cmd = ["/usr/bin/g++", "-std=c++17", "-O2", "-fPIC", "-Wall", "-Werror", "-shared", "-o", SO,
           "-I" + os.path.join(ROOT, "tests", "ros_shim"), "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(lib_dir, "host", "ros"),
           os.path.join(ROOT, "tests", "ros_shim", "adapter_harness.cpp"),
           os.path.join(lib_dir, "host", "ros", "b200_global_planner.cpp"),
           "-L" + lib_dir, "-lpp_b200", "-Wl,-rpath," + lib_dir]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return SO


def test_adapter_compiles_against_the_plugin_interface():
    so = build_adapter()
    syms = subprocess.run(["nm", "-D", "--defined-only", so], capture_output=True, text=True).stdout
    assert "adapter_make_plan" in syms
    # the plugin class overrides both pure virtuals of nav_core::BaseGlobalPlanner (it is instantiated in the harness)
    assert "B200GlobalPlanner" in subprocess.run(["nm", "-DC", so], capture_output=True, text=True).stdout


@pytest.mark.gpu
@pytest.mark.parametrize("motion,start,goal", [("ReedsShepp", (2.0, 2.0, 0.3), (17.5, 17.0, 1.2)),
                                               ("Dubins", (2.004, 2.009, 0.3), (12.0, 15.5, -2.0))])
def test_make_plan_equals_pp_plan(handle, motion, start, goal):
    from path_planner_b200.synth import make_costmap
    L = C.CDLL(build_adapter())
    pockets = [(int(start[0] / RES), int(start[1] / RES), 25), (int(goal[0] / RES), int(goal[1] / RES), 25)]
    cost = make_costmap(400, 400, seed=3, coverage=0.06, unknown_frac=0.0, pockets=pockets)
    boundary = np.array([0.64, 0.46], np.float32)
    s, g = np.array(start, np.float64), np.array(goal, np.float64)
    plan = np.zeros((65536, 3), np.float64)
    n, it = C.c_int(0), C.c_int(0)
    frame = C.create_string_buffer(32)
    ok = L.adapter_make_plan(cost.ctypes.data_as(C.c_void_p), 400, 400, C.c_double(RES), C.c_double(0.0), C.c_double(0.0),
                             boundary.ctypes.data_as(C.c_void_p), 2, 1, motion.encode(), 100000,
                             s.ctypes.data_as(C.c_void_p), g.ctypes.data_as(C.c_void_p), plan.ctypes.data_as(C.c_void_p),
                             65536, C.byref(n), C.byref(it), frame, 32)
    assert ok == 1 and frame.value == b"map"
    # the same request through the C ABI directly; poses truncated to centimetres like path_planner_ros.cpp:121-127
    import math
    cm = lambda v: np.float32(int(v * 100)) / np.float32(100.0)

    def yaw_through_quaternion(yaw):  # what tf::getYaw returns for the quaternion the harness builds
        z, w = math.sin(0.5 * yaw), math.cos(0.5 * yaw)
        return math.atan2(2.0 * (w * z), 1.0 - 2.0 * (z * z))

    as_sent = lambda p: (cm(p[0]), cm(p[1]), cm(yaw_through_quaternion(p[2])))
    handle.costmap_upload(cost)
    handle.footprint_from_boundary(RES, [0.64, 0.46], True)
    r = handle.plan(RES, (0.0, 0.0), as_sent(start), as_sent(goal), {"motion_type": 2 if motion == "ReedsShepp" else 1})
    assert r["ok"] and r["iterations"] == it.value and len(r["path"]) == n.value
    want = r["path"][::-1]  # the adapter reverses to start -> goal (path_planner_ros.cpp:140)
    got = plan[:n.value]
    assert np.array_equal(got[:, :2].astype(np.float32), want[:, :2])
    dyaw = np.angle(np.exp(1j * (got[:, 2] - want[:, 2].astype(np.float64))))  # through a quaternion and back
    assert np.abs(dyaw).max() < 1e-6
