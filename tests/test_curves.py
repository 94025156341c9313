"""CPU: the assumption path_planner_b200/host/curves.h makes about the host libm -- sin is odd and cos is even
bit for bit -- which lets one sin/cos pair serve the four symmetries of a Reeds-Shepp query."""
import ctypes

import numpy as np


def test_host_libm_sin_is_odd_and_cos_even_bitwise():
    libm = ctypes.CDLL("libm.so.6")
    for f in (libm.sin, libm.cos):
        f.restype = ctypes.c_double
        f.argtypes = [ctypes.c_double]
    rng = np.random.default_rng(3)
    xs = np.concatenate([rng.uniform(-7, 7, 40000), rng.uniform(-1e-3, 1e-3, 2000), rng.uniform(-400, 400, 8000),
                         np.array([0.0, np.pi, np.pi / 2, 1e-300, 3.0, 6.283185307179586])])
    for x in xs:
        x = float(x)
        assert np.float64(libm.sin(-x)).tobytes() == np.float64(-libm.sin(x)).tobytes()
        assert np.float64(libm.cos(-x)).tobytes() == np.float64(libm.cos(x)).tobytes()
