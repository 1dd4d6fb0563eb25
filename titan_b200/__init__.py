"""titan_b200 -- B200-native mass-spring step engine behind Titan's API.

Python here is plumbing only (ctypes onto the C ABI, build helper, bench/test drivers); the product is
libtitan_b200.so (CUDA kernels + C ABI, titan_b200/csrc) and libTitan.so (the C++ `titan::Simulation`
facade, titan_b200/host + include/Titan).
"""
from .capi import (Engine, Group, TitanError, Plane, Ball, Local, device_count, lib, LIB_PATH, EXPORTS,  # noqa: F401
                   MATH_FAST, MATH_STRICT, NO_GRAPH, KERNEL_CSR, INTEGRATOR_VERLET, NO_RESIDENT, INTEGRATOR_RK2)
from .build import build_all  # noqa: F401

__version__ = "0.1.0"
