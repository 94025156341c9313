"""path_planner_b200 -- B200-native (sm_100a) implementation of the costmap-parallel hot path of
liangheming/path_planner. The product is libpp_b200.so (C ABI, include/pp_b200.h) plus the C++
host classes in path_planner_b200/host/; `capi` is the ctypes binding the tests and bench use."""
from . import capi  # noqa: F401

__all__ = ["capi"]
