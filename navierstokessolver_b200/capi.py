"""ctypes binding of include/nsgpu.h (the drop-in boundary of the hot path).

Every call goes to libnsgpu.so; a missing library or a missing CUDA device raises (no fallback)."""
import ctypes as C
import os
import numpy as np

U, V, PHI, RHS_U, RHS_V, RHS_PHI, CONV_U0, CONV_V0, GRADPHI_X, GRADPHI_Y, USTAR, VSTAR = range(12)

_HERE = os.path.dirname(os.path.abspath(__file__))
lib_path = os.environ.get("NSGPU_LIB") or os.path.join(_HERE, "libnsgpu.so")   # NSGPU_LIB: another build of the same library (kernel experiments)


class NsGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("nsgpu error %d: %s" % (code, msg))
        self.code = code


class _Grid(C.Structure):
    _fields_ = [("nx", C.c_int32), ("ny", C.c_int32), ("n_cells", C.c_int32),
                ("hx", C.POINTER(C.c_double)), ("hy", C.POINTER(C.c_double)),
                ("cell_id", C.POINTER(C.c_int32)), ("cell_edges", C.POINTER(C.c_int32)),
                ("n_edges", C.c_int32),
                ("edge_nx", C.POINTER(C.c_int32)), ("edge_ny", C.POINTER(C.c_int32)),
                ("edge_bc_type", C.POINTER(C.c_int32)), ("edge_bc_info", C.POINTER(C.c_double))]


class _Params(C.Structure):
    _fields_ = [("dt", C.c_double), ("re", C.c_double), ("device", C.c_int32), ("rank", C.c_int32),
                ("world", C.c_int32), ("nccl_id", C.c_void_p), ("rtol_u", C.c_double), ("rtol_phi", C.c_double),
                ("max_cycles", C.c_int32), ("nu_pre", C.c_int32), ("nu_post", C.c_int32), ("solver", C.c_int32),
                ("use_graph", C.c_int32), ("max_levels", C.c_int32), ("vel_solver", C.c_int32), ("reserved", C.c_int32 * 6)]


class HostState(C.Structure):
    """nsgpu_host_state (include/nsgpu.h)."""
    _fields_ = [(n, C.POINTER(C.c_double)) for n in ("u", "v", "phi", "conv_u0", "conv_v0", "gradphi_x", "gradphi_y")]


class StepInfo(C.Structure):
    _fields_ = [("umin", C.c_double), ("umax", C.c_double), ("vmin", C.c_double), ("vmax", C.c_double),
                ("it_u", C.c_int32), ("it_v", C.c_int32), ("it_phi", C.c_int32),
                ("res_u", C.c_double), ("res_v", C.c_double), ("res_phi", C.c_double),
                ("fine_sweeps", C.c_double), ("ms_total", C.c_float), ("kernel_launches", C.c_int32),
                ("status", C.c_int32), ("flags", C.c_int32), ("krylov_restarts", C.c_int32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


def _load():
    if not os.path.exists(lib_path):
        raise ImportError("libnsgpu.so is not built (run `make` or __graft_entry__.build()); "
                          "navierstokessolver_b200 has no CPU fallback")
    # libnsgpu.so needs "libnccl.so.2".  torch bundles a newer NCCL under the same soname; load that copy first so
    # that a later `import torch` in the same process (bench.py, torch.distributed plumbing) finds its own symbols.
    import sys
    for d in sys.path:
        cand = os.path.join(d, "nvidia", "nccl", "lib", "libnccl.so.2")
        if os.path.exists(cand):
            try:
                C.CDLL(cand, mode=C.RTLD_GLOBAL)
                break
            except OSError:
                pass
    L = C.CDLL(lib_path)
    P = C.c_void_p
    dp = C.POINTER(C.c_double)
    L.nsgpu_nccl_unique_id.argtypes = [C.c_void_p]
    L.nsgpu_create.argtypes = [C.POINTER(_Grid), C.POINTER(_Params), C.POINTER(P)]
    L.nsgpu_destroy.argtypes = [P]
    L.nsgpu_destroy.restype = None
    L.nsgpu_last_error.argtypes = [P]
    L.nsgpu_last_error.restype = C.c_char_p
    L.nsgpu_local_range.argtypes = [P, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.nsgpu_step.argtypes = [P, C.c_int32, C.POINTER(StepInfo)]
    L.nsgpu_step_host.argtypes = [P, C.POINTER(HostState), C.POINTER(StepInfo)]
    L.nsgpu_phase_ms.argtypes = [P, C.POINTER(C.c_float)]
    L.nsgpu_rhs_build.argtypes = [P]
    L.nsgpu_solve_velocity.argtypes = [P, C.POINTER(StepInfo)]
    L.nsgpu_div_rhs.argtypes = [P, C.c_int32]
    L.nsgpu_solve_phi.argtypes = [P, C.POINTER(StepInfo)]
    L.nsgpu_correct.argtypes = [P, C.POINTER(StepInfo)]
    L.nsgpu_poisson_sweeps.argtypes = [P, C.c_int32, dp, C.POINTER(C.c_float)]
    L.nsgpu_apply_operator.argtypes = [P, C.c_int32, dp, dp]
    L.nsgpu_time_passes.argtypes = [P, C.c_int32, C.c_int32, C.POINTER(C.c_float)]
    L.nsgpu_pressure.argtypes = [P, dp, dp]
    L.nsgpu_export_begin.argtypes = [P, C.c_int64]
    L.nsgpu_export_wait.argtypes = [P] + [C.POINTER(dp)] * 5 + [C.POINTER(C.c_int64)]
    L.nsgpu_export_last_ms.argtypes = [P, C.POINTER(C.c_float)]
    L.nsgpu_minmax_noalloc.argtypes = [C.c_int32, dp, dp, C.c_int64, dp]
    L.nsgpu_noalloc_max_cells.restype = C.c_longlong
    L.nsgpu_get_array.argtypes = [P, C.c_int32, dp]
    L.nsgpu_set_array.argtypes = [P, C.c_int32, dp]
    L.nsgpu_get_fields.argtypes = [P, dp, dp, dp]
    L.nsgpu_set_fields.argtypes = [P, dp, dp, dp]
    L.nsgpu_num_levels.argtypes = [P]
    L.nsgpu_kernel_launches.argtypes = [P]
    L.nsgpu_kernel_launches.restype = C.c_int64
    L.nsgpu_synchronize.argtypes = [P]
    L.nsgpu_checkpoint_save.argtypes = [P, C.c_char_p, C.c_int64]
    L.nsgpu_checkpoint_load.argtypes = [P, C.c_char_p, C.POINTER(C.c_int64)]
    return L


lib = _load()

EXPORTED = ["nsgpu_nccl_unique_id", "nsgpu_create", "nsgpu_destroy", "nsgpu_last_error", "nsgpu_local_range",
            "nsgpu_step", "nsgpu_rhs_build", "nsgpu_solve_velocity", "nsgpu_div_rhs", "nsgpu_solve_phi",
            "nsgpu_correct", "nsgpu_poisson_sweeps", "nsgpu_apply_operator", "nsgpu_pressure", "nsgpu_get_array",
            "nsgpu_set_array", "nsgpu_get_fields", "nsgpu_set_fields", "nsgpu_num_levels", "nsgpu_kernel_launches",
            "nsgpu_synchronize", "nsgpu_phase_ms", "nsgpu_step_host", "nsgpu_checkpoint_save", "nsgpu_checkpoint_load",
            "nsgpu_export_begin", "nsgpu_export_wait", "nsgpu_time_passes", "nsgpu_export_last_ms",
            "nsgpu_minmax_noalloc", "nsgpu_noalloc_max_cells"]


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


class GridDesc:
    """Flattened copy of the reference's Grid public data (Grid.h:58-67), numpy arrays."""

    def __init__(self, hx, hy, edge_nx, edge_ny, edge_bc_type, edge_bc_info, cell_id=None, cell_edges=None, n_cells=None):
        self.hx = np.ascontiguousarray(hx, dtype=np.float64)
        self.hy = np.ascontiguousarray(hy, dtype=np.float64)
        self.nx, self.ny = len(self.hx), len(self.hy)
        self.edge_nx = np.ascontiguousarray(edge_nx, dtype=np.int32)
        self.edge_ny = np.ascontiguousarray(edge_ny, dtype=np.int32)
        self.edge_bc_type = np.ascontiguousarray(edge_bc_type, dtype=np.int32)
        self.edge_bc_info = np.ascontiguousarray(edge_bc_info, dtype=np.float64)
        self.cell_id = None if cell_id is None else np.ascontiguousarray(cell_id, dtype=np.int32).reshape(-1)
        self.cell_edges = None if cell_edges is None else np.ascontiguousarray(cell_edges, dtype=np.int32).reshape(-1)
        if n_cells is None:
            n_cells = self.nx * self.ny if self.cell_id is None else int((self.cell_id >= 0).sum())
        self.n_cells = int(n_cells)

    @staticmethod
    def cavity(n, lid=1.0, nx=None, lx=1.0, ly=1.0):
        """Uniform rectangle, all walls, moving top lid: edges in the reference's order left, top, right, bottom."""
        nx = n if nx is None else nx
        return GridDesc(np.full(nx, lx / nx), np.full(n, ly / n), [-1, 0, 1, 0], [0, 1, 0, -1], [2, 2, 2, 2], [0.0, lid, 0.0, 0.0])

    def c_struct(self):
        g = _Grid()
        g.nx, g.ny, g.n_cells = self.nx, self.ny, self.n_cells
        g.hx, g.hy = _dp(self.hx), _dp(self.hy)
        g.cell_id = _ip(self.cell_id) if self.cell_id is not None else None
        g.cell_edges = _ip(self.cell_edges) if self.cell_edges is not None else None
        g.n_edges = len(self.edge_nx)
        g.edge_nx, g.edge_ny, g.edge_bc_type = _ip(self.edge_nx), _ip(self.edge_ny), _ip(self.edge_bc_type)
        g.edge_bc_info = _dp(self.edge_bc_info)
        return g


def minmax_noalloc(u, v, device=0):
    """(umin, umax, vmin, vmax) through nsgpu_minmax_noalloc: no allocation on the device, scratch in a static device array."""
    u = np.ascontiguousarray(u, dtype=np.float64)
    v = np.ascontiguousarray(v, dtype=np.float64)
    out = np.empty(4)
    rc = lib.nsgpu_minmax_noalloc(device, _dp(u), _dp(v), u.size, _dp(out))
    if rc:
        raise NsGpuError(rc, lib.nsgpu_last_error(None).decode())
    return tuple(float(x) for x in out)


def nccl_unique_id():
    buf = C.create_string_buffer(128)
    rc = lib.nsgpu_nccl_unique_id(buf)
    if rc:
        raise NsGpuError(rc, "ncclGetUniqueId failed")
    return buf.raw


class GpuFluidSolver:
    """One nsgpu_ctx: the device-resident state of a FluidSolver (FS.h:20-59)."""

    def __init__(self, grid, dt, re, device=0, rank=0, world=1, nccl_id=None, rtol_u=0.0, rtol_phi=0.0, max_cycles=0,
                 nu_pre=0, nu_post=0, solver=0, use_graph=0, max_levels=0, vel_solver=0):
        self.grid = grid
        self._g = grid.c_struct()
        p = _Params()
        p.dt, p.re, p.device, p.rank, p.world = dt, re, device, rank, world
        self._nccl = C.create_string_buffer(nccl_id, 128) if nccl_id is not None else None
        p.nccl_id = C.cast(self._nccl, C.c_void_p) if self._nccl is not None else None
        p.rtol_u, p.rtol_phi, p.max_cycles = rtol_u, rtol_phi, max_cycles
        p.nu_pre, p.nu_post, p.solver, p.use_graph, p.max_levels = nu_pre, nu_post, solver, use_graph, max_levels
        p.vel_solver = vel_solver
        self._h = C.c_void_p()
        rc = lib.nsgpu_create(C.byref(self._g), C.byref(p), C.byref(self._h))
        if rc:
            raise NsGpuError(rc, lib.nsgpu_last_error(None).decode())
        i0, i1, id0, n = C.c_int32(), C.c_int32(), C.c_int64(), C.c_int64()
        lib.nsgpu_local_range(self._h, C.byref(i0), C.byref(i1), C.byref(id0), C.byref(n))
        self.i0, self.i1, self.id0, self.n_local = i0.value, i1.value, id0.value, n.value

    def close(self):
        if getattr(self, "_h", None) and lib is not None:   # `lib` is already gone at interpreter shutdown
            lib.nsgpu_destroy(self._h)
            self._h = None

    __del__ = close

    def _chk(self, rc, soft=()):
        if rc and rc not in soft:
            raise NsGpuError(rc, lib.nsgpu_last_error(self._h).decode())
        return rc

    def get(self, which):
        out = np.empty(self.n_local, dtype=np.float64)
        self._chk(lib.nsgpu_get_array(self._h, which, _dp(out)))
        return out

    def set(self, which, a):
        a = np.ascontiguousarray(a, dtype=np.float64)
        assert a.size == self.n_local
        self._chk(lib.nsgpu_set_array(self._h, which, _dp(a)))

    def step(self, n=1):
        info = StepInfo()
        self._chk(lib.nsgpu_step(self._h, n, C.byref(info)), soft=(-5,))
        return info

    def step_host(self, state):
        """One time step with the whole state in host arrays (nsgpu_step_host): `state` maps U, V, PHI, CONV_U0, CONV_V0,
        GRADPHI_X, GRADPHI_Y to contiguous float64 arrays of this rank's id range; they are overwritten with the new state.
        Without the two GRADPHI entries divPhi is rebuilt on the device from PHI and not shipped (five arrays each way).
        Page-locked arrays make the copies overlap the phases."""
        with_grad = GRADPHI_X in state
        io = HostState(*[_dp(state[k]) for k in (U, V, PHI, CONV_U0, CONV_V0)],
                       _dp(state[GRADPHI_X]) if with_grad else None, _dp(state[GRADPHI_Y]) if with_grad else None)
        info = StepInfo()
        self._chk(lib.nsgpu_step_host(self._h, C.byref(io), C.byref(info)), soft=(-5,))
        return info

    def checkpoint_save(self, path, step=0):
        """Complete state of the time loop (7 arrays of this rank's id range) + the caller's step counter -> file (nsgpu_checkpoint_save)."""
        self._chk(lib.nsgpu_checkpoint_save(self._h, str(path).encode(), int(step)))

    def checkpoint_load(self, path):
        """Restores the state written by checkpoint_save on the same grid and slab decomposition; returns the stored step counter."""
        step = C.c_int64(0)
        self._chk(lib.nsgpu_checkpoint_load(self._h, str(path).encode(), C.byref(step)))
        return int(step.value)

    def rhs_build(self):
        self._chk(lib.nsgpu_rhs_build(self._h))

    def solve_velocity(self):
        info = StepInfo()
        self._chk(lib.nsgpu_solve_velocity(self._h, C.byref(info)), soft=(-5,))
        return info

    def div_rhs(self, remove_mean=True):
        self._chk(lib.nsgpu_div_rhs(self._h, 1 if remove_mean else 0))

    def solve_phi(self):
        info = StepInfo()
        self._chk(lib.nsgpu_solve_phi(self._h, C.byref(info)), soft=(-5,))
        return info

    def correct(self):
        info = StepInfo()
        self._chk(lib.nsgpu_correct(self._h, C.byref(info)))
        return info

    def poisson_sweeps(self, n, want_residual=True):
        r, ms = C.c_double(0.0), C.c_float(0.0)
        self._chk(lib.nsgpu_poisson_sweeps(self._h, n, C.byref(r) if want_residual else None, C.byref(ms)))
        return r.value, ms.value

    def time_passes(self, kind, n):
        """device ms of n back-to-back fine-level passes of `kind` (nsgpu_time_passes)"""
        ms = C.c_float(0.0)
        self._chk(lib.nsgpu_time_passes(self._h, kind, n, C.byref(ms)))
        return ms.value

    def apply_operator(self, op, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty(self.n_local, dtype=np.float64)
        self._chk(lib.nsgpu_apply_operator(self._h, op, _dp(x), _dp(y)))
        return y

    def pressure(self):
        pr, lap = np.empty(self.n_local), np.empty(self.n_local)
        self._chk(lib.nsgpu_pressure(self._h, _dp(pr), _dp(lap)))
        return pr, lap

    def export_begin(self, tag=0):
        """Snapshot of u, v, phi, Pr, lapPhi on the device + asynchronous copy home (nsgpu_export_begin); returns at once."""
        self._chk(lib.nsgpu_export_begin(self._h, int(tag)))

    def export_wait(self):
        """Oldest pending snapshot: (tag, {"u", "v", "phi", "pr", "lap"}) as copies of the context's pinned arrays."""
        ptr = [C.POINTER(C.c_double)() for _ in range(5)]
        tag = C.c_int64(0)
        self._chk(lib.nsgpu_export_wait(self._h, *[C.byref(q) for q in ptr], C.byref(tag)))
        names = ("u", "v", "phi", "pr", "lap")
        return int(tag.value), {k: np.ctypeslib.as_array(q, shape=(self.n_local,)).copy() for k, q in zip(names, ptr)}

    @property
    def num_levels(self):
        return lib.nsgpu_num_levels(self._h)

    @property
    def kernel_launches(self):
        return lib.nsgpu_kernel_launches(self._h)

    def phase_ms(self):
        """device ms of (rhs build, velocity solves, div + mean, phi solve, correct + min/max) in the last step run"""
        out = (C.c_float * 5)()
        self._chk(lib.nsgpu_phase_ms(self._h, out))
        return [float(x) for x in out]

    def synchronize(self):
        self._chk(lib.nsgpu_synchronize(self._h))
