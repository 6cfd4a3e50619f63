"""navierstokessolver_b200 -- B200-native time-step loop of shivams15/navierstokessolver.

The package is a thin host-side mirror of the reference's FluidSolver/Grid interface over the C ABI of
``libnsgpu.so`` (include/nsgpu.h).  All arithmetic runs in hand-written sm_100a kernels; there is no CPU
fallback: importing :mod:`navierstokessolver_b200.capi` fails loudly when the library is missing.
"""
from .capi import (  # noqa: F401
    GpuFluidSolver, GridDesc, NsGpuError, lib, lib_path,
    U, V, PHI, RHS_U, RHS_V, RHS_PHI, CONV_U0, CONV_V0, GRADPHI_X, GRADPHI_Y, USTAR, VSTAR,
)
from .host import HostGrid, HostFluidSolver, load_case, minmode  # noqa: F401,E402
