"""Python mirror of the host classes (navierstokessolver_b200/host/Grid.h, FluidSolver.h) through the nshost_* C wrappers
compiled into libnsgpu.so.  HostGrid parses a grid file exactly like the C++ Grid; HostFluidSolver is FluidSolver."""
import ctypes as C
import os
import numpy as np

from .capi import lib, GridDesc

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)

lib.nshost_grid_create.restype = C.c_void_p
lib.nshost_grid_create.argtypes = [C.c_char_p, C.c_int]
lib.nshost_log.restype = C.c_char_p
lib.nshost_log.argtypes = [C.c_void_p]
for _f in ("nshost_grid_setup", "nshost_n", "nshost_nx", "nshost_ny", "nshost_n_edges"):
    getattr(lib, _f).argtypes = [C.c_void_p]
lib.nshost_grid_arrays.argtypes = [C.c_void_p, _dp, _dp, _ip, _ip, _dp, _dp]
lib.nshost_edges.argtypes = [C.c_void_p, _ip, _ip, _ip, _dp, _dp]
lib.nshost_solver_create.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
lib.nshost_solve.argtypes = [C.c_void_p]
lib.nshost_solve.restype = None
lib.nshost_device.argtypes = [C.c_void_p]
lib.nshost_device.restype = C.c_void_p
lib.nshost_destroy.argtypes = [C.c_void_p]
lib.nshost_destroy.restype = None
lib.nshost_minmode.argtypes = [C.c_double, C.c_double]
lib.nshost_minmode.restype = C.c_double


def minmode(a, b):
    return lib.nshost_minmode(a, b)


class HostGrid:
    """Grid(char* fname) (Grid.h).  Writes CellCenters.csv into the CWD like the reference (Grid.cpp:15)."""

    def __init__(self, gridfile, quiet=True):
        self.h = lib.nshost_grid_create(os.fsencode(gridfile), 1 if quiet else 0)
        self.log = lib.nshost_log(self.h).decode()
        self.setup = bool(lib.nshost_grid_setup(self.h))
        self.fs_setup = False
        if not self.setup:
            return
        self.N, self.nx, self.ny = lib.nshost_n(self.h), lib.nshost_nx(self.h), lib.nshost_ny(self.h)
        self.hx, self.hy = np.empty(self.nx), np.empty(self.ny)
        self.cell_id = np.empty(self.nx * self.ny, dtype=np.int32)
        self.cell_edges = np.empty(self.nx * self.ny * 4, dtype=np.int32)
        self.xc, self.yc = np.empty(self.nx * self.ny), np.empty(self.nx * self.ny)
        lib.nshost_grid_arrays(self.h, self.hx.ctypes.data_as(_dp), self.hy.ctypes.data_as(_dp), self.cell_id.ctypes.data_as(_ip),
                               self.cell_edges.ctypes.data_as(_ip), self.xc.ctypes.data_as(_dp), self.yc.ctypes.data_as(_dp))
        self._read_edges()

    def _read_edges(self):
        ne = lib.nshost_n_edges(self.h)
        self.edge_nx, self.edge_ny = np.empty(ne, dtype=np.int32), np.empty(ne, dtype=np.int32)
        self.edge_type, self.edge_info, self.edge_loc = np.empty(ne, dtype=np.int32), np.empty(ne), np.empty(3 * ne)
        lib.nshost_edges(self.h, self.edge_nx.ctypes.data_as(_ip), self.edge_ny.ctypes.data_as(_ip), self.edge_type.ctypes.data_as(_ip),
                         self.edge_info.ctypes.data_as(_dp), self.edge_loc.ctypes.data_as(_dp))

    def desc(self):
        return GridDesc(self.hx, self.hy, self.edge_nx, self.edge_ny, self.edge_type, self.edge_info, self.cell_id, self.cell_edges, self.N)

    def close(self):
        if getattr(self, "h", None):
            lib.nshost_destroy(self.h)
            self.h = None

    __del__ = close


class HostFluidSolver:
    """FluidSolver(char* fname, Grid* grid) + Solve() (FluidSolver.h); needs a CUDA device."""

    def __init__(self, simfile, grid, quiet=True):
        self.grid = grid
        self.setup = bool(lib.nshost_solver_create(grid.h, os.fsencode(simfile), 1 if quiet else 0))
        self.log = lib.nshost_log(grid.h).decode()
        grid._read_edges()   # the parser wrote bcType / bcInfo into the grid's edges (FS.cpp:651)

    def Solve(self):
        lib.nshost_solve(self.grid.h)

    def device(self):
        return lib.nshost_device(self.grid.h)


def load_case(gridfile, quiet=True):
    return HostGrid(gridfile, quiet)
