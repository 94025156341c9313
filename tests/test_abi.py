"""CPU: the C-ABI library loads and exports every symbol include/pp_b200.h declares. No compute
call is made here (there is no GPU in the build container and no CPU path in the library)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT
from path_planner_b200 import capi


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "pp_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pp_[a-z0-9_]+)\s*\(", text)))


def test_library_is_built():
    assert os.path.exists(capi.LIB_PATH), "run __graft_entry__.build() first"


def test_exports_every_declared_symbol():
    lib = ctypes.CDLL(capi.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_binding_covers_header():
    assert sorted(capi.SYMBOLS) == [n for n in declared_symbols() if n in capi.SYMBOLS]
    not_bound = [n for n in declared_symbols() if n not in capi.SYMBOLS]
    assert not not_bound, not_bound


def test_version_string():
    assert b"sm_100a" in capi.load().pp_version()


def test_no_oracle_in_product():
    """The product must never reach into oracle/ (tests, smoke and bench's CPU leg only)."""
    bad = []
    for base, _, files in os.walk(os.path.join(ROOT, "path_planner_b200")):
        if "build" in base:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".hpp")):
                s = open(os.path.join(base, f), errors="ignore").read()
                if re.search(r"\bimport oracle\b|from oracle\b|liboracle_port|libpp_ref|oracle/", s):
                    bad.append(f)
    assert not bad, bad


def test_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(capi.PPError):
        capi.Handle(0)


def test_window_origin_is_host_arithmetic():
    """pp_window_origin_cell is plain host arithmetic (Costmap2D::updateOrigin + the static layer's worldToMap): it
    runs without a GPU and equals the oracle's restatement."""
    import numpy as np
    import oracle
    L = capi.load()
    rng = np.random.default_rng(3)
    cases = [(5.0, 10.0, 0.05, 0.0), (12.37, 10.0, 0.05, 0.0), (0.3, 10.0, 0.05, 0.0), (12.37, 10.0, 0.05, -10.0), (-3.21, 6.0, 0.1, -50.0)]
    cases += [(float(rng.uniform(-40, 40)), 10.0, 0.05, float(rng.uniform(-20, 0))) for _ in range(200)]
    for robot, size, res, origin in cases:
        assert L.pp_window_origin_cell(robot, size, res, origin) == oracle.Port.window_origin_cell(robot, size, res, origin), (robot, size, res, origin)
