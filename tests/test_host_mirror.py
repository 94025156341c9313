"""The C++ host mirror of the follower-side interface (path_planner_b200/host/obstacle_index.h): compiled
with -Wall -Werror on the CPU; on the GPU it returns what the C ABI returns through the ctypes harness."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

SO = os.path.join(ROOT, "tests", "host_harness", "libpp_host_harness_test.so")


def build_harness():
    lib_dir = os.path.join(ROOT, "path_planner_b200")
    cmd = ["/usr/bin/g++", "-std=c++17", "-O2", "-fPIC", "-Wall", "-Werror", "-shared", "-o", SO,
           "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(lib_dir, "host"),
           os.path.join(ROOT, "tests", "host_harness", "obstacle_index_harness.cpp"),
           "-L" + lib_dir, "-lpp_b200", "-Wl,-rpath," + lib_dir]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return SO


def test_obstacle_index_header_compiles():
    so = build_harness()
    assert "harness_obstacle_index" in subprocess.run(["nm", "-D", "--defined-only", so], capture_output=True, text=True).stdout


@pytest.mark.gpu
def test_obstacle_index_equals_the_c_abi(handle):
    L = C.CDLL(build_harness())
    rng = np.random.default_rng(23)
    pts = np.ascontiguousarray(rng.uniform(-3, 3, (4000, 2)))
    poses = np.ascontiguousarray(np.concatenate([rng.uniform(-2.5, 2.5, (60, 2)), rng.uniform(-3.2, 3.2, (60, 1))], axis=1))
    robot = np.array([0.72, 0.50])
    Q, K = len(poses), 10
    idx, d2, cnt = np.zeros((Q, K), np.uint32), np.zeros((Q, K)), np.zeros(Q, np.uint32)
    err, jac, radii = np.zeros((Q, K, 2)), np.zeros((Q, K, 2, 3)), np.zeros(2)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    ok = L.harness_obstacle_index(0, p(pts), len(pts), p(poses), Q, C.c_double(0.9574), K, p(robot), 2, p(idx), p(d2), p(cnt),
                                  p(err), p(jac), p(radii))
    assert ok == 1
    i0, d0, c0 = handle.radius_topk(pts, poses[:, :2], 0.9574, K)
    used = np.arange(K)[None, :] < c0[:, None]
    assert np.array_equal(cnt, c0) and np.array_equal(idx[used], i0[used]) and np.array_equal(d2[used], d0[used])
    assert tuple(radii) == handle.robot_model_set(robot)
    e0, j0 = handle.obstacle_edges(poses, K)
    assert np.array_equal(err[used], e0[used]) and np.array_equal(jac[used], j0[used])
