"""TEST INFRASTRUCTURE -- ctypes access to oracle/_ref/libnsref.so (the unmodified reference classes built against the
PETSc stand-in, see oracle/ref_driver.cpp) and to tests/emul/libnsemul.so (host compile of the device functions)."""
import ctypes as C
import os
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libnsref.so")
EMUL_SO = os.path.join(ROOT, "tests", "emul", "libnsemul.so")
CASES = os.path.join(ROOT, "cases")

U, V, PHI, RHS_U, RHS_V, RHS_PHI, CONV_U0, CONV_V0, GRADPHI_X, GRADPHI_Y, USTAR, VSTAR = range(12)

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def have_ref():
    return os.path.exists(REF_SO)


_ref = None


def ref_lib():
    global _ref
    if _ref is None:
        L = C.CDLL(REF_SO)
        L.nsref_create.restype = C.c_void_p
        L.nsref_create.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.nsref_grid_create.restype = C.c_void_p
        L.nsref_grid_create.argtypes = [C.c_char_p, C.c_int]
        for f in ("nsref_n", "nsref_nx", "nsref_ny", "nsref_n_edges", "nsref_grid_setup"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.nsref_grid_arrays.argtypes = [C.c_void_p, _dp, _dp, _ip, _ip, _dp, _dp]
        L.nsref_edges.argtypes = [C.c_void_p, _ip, _ip, _ip, _dp, _dp]
        L.nsref_param.argtypes = [C.c_void_p, C.c_int]
        L.nsref_param.restype = C.c_double
        L.nsref_get.argtypes = [C.c_void_p, C.c_int, _dp]
        L.nsref_set.argtypes = [C.c_void_p, C.c_int, _dp]
        for f in ("nsref_construct_rhs_v", "nsref_solve_velocity", "nsref_construct_rhs_phi", "nsref_nullspace_remove",
                  "nsref_solve_phi", "nsref_correct"):
            getattr(L, f).argtypes = [C.c_void_p]
            getattr(L, f).restype = None
        L.nsref_step.argtypes = [C.c_void_p, C.c_int, _dp]
        L.nsref_step.restype = None
        L.nsref_phase_seconds.argtypes = [C.c_void_p, _dp]
        L.nsref_mat_nnz.argtypes = [C.c_void_p, C.c_int]
        L.nsref_mat_csr.argtypes = [C.c_void_p, C.c_int, _ip, _ip, _dp]
        L.nsref_ksp_stats.argtypes = [C.c_void_p, C.c_int, _ip, _ip, _dp, _dp, _ip, _dp, C.POINTER(C.c_long), C.POINTER(C.c_long)]
        L.nsref_set_rtol.argtypes = [C.c_void_p, C.c_double]
        L.nsref_export.argtypes = [C.c_void_p, C.c_int]
        L.nsref_minmode.argtypes = [C.c_double, C.c_double]
        L.nsref_minmode.restype = C.c_double
        _ref = L
    return _ref


class RefSolver:
    """The reference's Grid + FluidSolver, phase by phase.  Writes CellCenters.csv into the CWD (Grid.cpp:15)."""

    def __init__(self, gridfile, simfile, quiet=True, workdir=None):
        L = ref_lib()
        cwd = os.getcwd()
        if workdir:
            os.chdir(workdir)
        try:
            self.h = L.nsref_create(gridfile.encode(), simfile.encode(), 1 if quiet else 0)
        finally:
            os.chdir(cwd)
        if not self.h:
            raise RuntimeError("reference setup failed for %s / %s" % (gridfile, simfile))
        self.L = L
        self.n, self.nx, self.ny, self.n_edges = L.nsref_n(self.h), L.nsref_nx(self.h), L.nsref_ny(self.h), L.nsref_n_edges(self.h)
        self.dt, self.final_time, self.re = (L.nsref_param(self.h, k) for k in range(3))
        self.hx, self.hy = np.empty(self.nx), np.empty(self.ny)
        self.cell_id = np.empty(self.nx * self.ny, dtype=np.int32)
        self.cell_edges = np.empty(self.nx * self.ny * 4, dtype=np.int32)
        self.xc, self.yc = np.empty(self.nx * self.ny), np.empty(self.nx * self.ny)
        L.nsref_grid_arrays(self.h, self.hx.ctypes.data_as(_dp), self.hy.ctypes.data_as(_dp), self.cell_id.ctypes.data_as(_ip),
                            self.cell_edges.ctypes.data_as(_ip), self.xc.ctypes.data_as(_dp), self.yc.ctypes.data_as(_dp))
        self.edge_nx, self.edge_ny = np.empty(self.n_edges, dtype=np.int32), np.empty(self.n_edges, dtype=np.int32)
        self.edge_type, self.edge_info = np.empty(self.n_edges, dtype=np.int32), np.empty(self.n_edges)
        L.nsref_edges(self.h, self.edge_nx.ctypes.data_as(_ip), self.edge_ny.ctypes.data_as(_ip), self.edge_type.ctypes.data_as(_ip),
                      self.edge_info.ctypes.data_as(_dp), None)

    def desc_kwargs(self):
        return dict(hx=self.hx, hy=self.hy, edge_nx=self.edge_nx, edge_ny=self.edge_ny, edge_bc_type=self.edge_type,
                    edge_bc_info=self.edge_info, cell_id=self.cell_id, cell_edges=self.cell_edges, n_cells=self.n)

    def get(self, which):
        out = np.empty(self.n)
        assert self.L.nsref_get(self.h, which, out.ctypes.data_as(_dp)) == 0
        return out

    def set(self, which, a):
        a = np.ascontiguousarray(a, dtype=np.float64)
        assert a.size == self.n and self.L.nsref_set(self.h, which, a.ctypes.data_as(_dp)) == 0

    def construct_rhs_v(self): self.L.nsref_construct_rhs_v(self.h)
    def solve_velocity(self): self.L.nsref_solve_velocity(self.h)
    def construct_rhs_phi(self): self.L.nsref_construct_rhs_phi(self.h)
    def nullspace_remove(self): self.L.nsref_nullspace_remove(self.h)
    def solve_phi(self): self.L.nsref_solve_phi(self.h)
    def correct(self): self.L.nsref_correct(self.h)
    def set_rtol(self, r): self.L.nsref_set_rtol(self.h, r)

    def step(self, n=1):
        mm = np.empty(4)
        self.L.nsref_step(self.h, n, mm.ctypes.data_as(_dp))
        return mm

    def phase_seconds(self):
        t = np.empty(5)
        self.L.nsref_phase_seconds(self.h, t.ctypes.data_as(_dp))
        return t

    def ksp_stats(self, which):
        its, mv, conv = C.c_int(), C.c_int(), C.c_int()
        r0, r, sec = C.c_double(), C.c_double(), C.c_double()
        solves, tot = C.c_long(), C.c_long()
        self.L.nsref_ksp_stats(self.h, which, C.byref(its), C.byref(mv), C.byref(r0), C.byref(r), C.byref(conv), C.byref(sec),
                               C.byref(solves), C.byref(tot))
        return dict(its=its.value, matvecs=mv.value, rnorm0=r0.value, rnorm=r.value, converged=conv.value, seconds=sec.value,
                    solves=solves.value, iterations_total=tot.value)

    def csr(self, which):
        import scipy.sparse as sp
        nnz = self.L.nsref_mat_nnz(self.h, which)
        rp, col, val = np.empty(self.n + 1, dtype=np.int32), np.empty(nnz, dtype=np.int32), np.empty(nnz)
        assert self.L.nsref_mat_csr(self.h, which, rp.ctypes.data_as(_ip), col.ctypes.data_as(_ip), val.ctypes.data_as(_dp)) == 0
        return sp.csr_matrix((val, col, rp), shape=(self.n, self.n))


# ------------------------------------------------------------------------------------------------------------------
class _Grid(C.Structure):
    _fields_ = [("nx", C.c_int32), ("ny", C.c_int32), ("n_cells", C.c_int32),
                ("hx", _dp), ("hy", _dp), ("cell_id", C.POINTER(C.c_int32)), ("cell_edges", C.POINTER(C.c_int32)),
                ("n_edges", C.c_int32), ("edge_nx", C.POINTER(C.c_int32)), ("edge_ny", C.POINTER(C.c_int32)),
                ("edge_bc_type", C.POINTER(C.c_int32)), ("edge_bc_info", _dp)]


_emul = None


def emul_lib():
    global _emul
    if _emul is None:
        L = C.CDLL(EMUL_SO)
        L.emul_create.restype = C.c_void_p
        L.emul_create.argtypes = [C.POINTER(_Grid), C.c_double, C.c_double, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int]
        L.emul_create_rank.restype = C.c_void_p
        L.emul_create_rank.argtypes = [C.POINTER(_Grid), C.c_double, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int]
        L.emul_plan_info.argtypes = [C.c_void_p, C.POINTER(C.c_long), _ip, _ip]
        L.emul_ld.argtypes = [C.c_void_p]
        L.emul_halo_pack.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, _dp]
        L.emul_halo_unpack.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, _dp]
        L.emul_phi_half_sweep.argtypes = [C.c_void_p, C.c_int]
        L.emul_destroy.argtypes = [C.c_void_p]
        L.emul_rect.argtypes = [C.c_void_p]
        L.emul_num_levels.argtypes = [C.c_void_p]
        L.emul_n_local.argtypes = [C.c_void_p]
        L.emul_n_local.restype = C.c_long
        L.emul_level_shape.argtypes = [C.c_void_p, C.c_int, _ip, _ip, _ip, _ip]
        L.emul_set.argtypes = [C.c_void_p, C.c_int, _dp]
        L.emul_get.argtypes = [C.c_void_p, C.c_int, _dp]
        L.emul_rhs_build.argtypes = [C.c_void_p]
        L.emul_div_rhs.argtypes = [C.c_void_p, C.c_int]
        L.emul_correct.argtypes = [C.c_void_p, _dp]
        L.emul_apply.argtypes = [C.c_void_p, C.c_int, _dp, _dp]
        L.emul_solve.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_int, _dp]
        L.emul_poisson_sweeps.argtypes = [C.c_void_p, C.c_int, _dp]
        L.emul_step.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_int, _dp, _ip]
        L.emul_step.restype = C.c_double
        L.emul_phi_ilu_w.argtypes = [C.c_void_p, _dp]
        L.emul_phi_ilu_w.restype = C.c_double
        _emul = L
    return _emul


class Emul:
    """Host execution of the device functions (tests only)."""

    def __init__(self, hx, hy, edge_nx, edge_ny, edge_bc_type, edge_bc_info, dt, re, cell_id=None, cell_edges=None, n_cells=None,
                 nu_pre=2, nu_post=2, max_levels=0, rank=0, world=1):
        L = emul_lib()
        self._keep = [np.ascontiguousarray(hx, dtype=np.float64), np.ascontiguousarray(hy, dtype=np.float64),
                      np.ascontiguousarray(edge_nx, dtype=np.int32), np.ascontiguousarray(edge_ny, dtype=np.int32),
                      np.ascontiguousarray(edge_bc_type, dtype=np.int32), np.ascontiguousarray(edge_bc_info, dtype=np.float64)]
        g = _Grid()
        g.nx, g.ny = len(self._keep[0]), len(self._keep[1])
        g.hx, g.hy = self._keep[0].ctypes.data_as(_dp), self._keep[1].ctypes.data_as(_dp)
        g.n_edges = len(self._keep[2])
        i32 = C.POINTER(C.c_int32)
        g.edge_nx, g.edge_ny, g.edge_bc_type = (self._keep[k].ctypes.data_as(i32) for k in (2, 3, 4))
        g.edge_bc_info = self._keep[5].ctypes.data_as(_dp)
        if cell_id is not None:
            ci, ce = np.ascontiguousarray(cell_id, dtype=np.int32), np.ascontiguousarray(cell_edges, dtype=np.int32)
            self._keep += [ci, ce]
            g.cell_id, g.cell_edges = ci.ctypes.data_as(i32), ce.ctypes.data_as(i32)
            g.n_cells = int((ci >= 0).sum()) if n_cells is None else n_cells
        else:
            g.n_cells = g.nx * g.ny
        err = C.create_string_buffer(512)
        self.h = L.emul_create_rank(C.byref(g), dt, re, nu_pre, nu_post, max_levels, rank, world, err, 512)
        if not self.h:
            raise RuntimeError(err.value.decode())
        self.L = L
        self.n = L.emul_n_local(self.h)
        self.rect = bool(L.emul_rect(self.h))
        self.num_levels = L.emul_num_levels(self.h)

    def plan_info(self):
        """Slab plan of this rank: dict(i0, i1, id0, n_local, n_levels, rep_start, lnx, lgi0)."""
        out = (C.c_long * 6)()
        lnx, lgi0 = (C.c_int * 64)(), (C.c_int * 64)()
        self.L.emul_plan_info(self.h, out, lnx, lgi0)
        n = int(out[4])
        return dict(i0=int(out[0]), i1=int(out[1]), id0=int(out[2]), n_local=int(out[3]), n_levels=n, rep_start=int(out[5]),
                    lnx=[int(lnx[k]) for k in range(n)], lgi0=[int(lgi0[k]) for k in range(n)])

    def halo_pack(self, which, side, width):
        out = np.empty(width * self.L.emul_ld(self.h))
        self.L.emul_halo_pack(self.h, which, side, width, out.ctypes.data_as(_dp))
        return out

    def halo_unpack(self, which, side, width, data):
        data = np.ascontiguousarray(data, dtype=np.float64)
        assert data.size == width * self.L.emul_ld(self.h)
        self.L.emul_halo_unpack(self.h, which, side, width, data.ctypes.data_as(_dp))

    def phi_half_sweep(self, colour):
        self.L.emul_phi_half_sweep(self.h, colour)

    def level_shapes(self):
        out = []
        for l in range(self.num_levels):
            a, b, c, d = C.c_int(), C.c_int(), C.c_int(), C.c_int()
            self.L.emul_level_shape(self.h, l, C.byref(a), C.byref(b), C.byref(c), C.byref(d))
            out.append((a.value, b.value, c.value, d.value))
        return out

    def get(self, which):
        out = np.zeros(self.n)
        self.L.emul_get(self.h, which, out.ctypes.data_as(_dp))
        return out

    def set(self, which, a):
        a = np.ascontiguousarray(a, dtype=np.float64)
        assert a.size == self.n
        self.L.emul_set(self.h, which, a.ctypes.data_as(_dp))

    def rhs_build(self): self.L.emul_rhs_build(self.h)
    def div_rhs(self, remove_mean=True): self.L.emul_div_rhs(self.h, 1 if remove_mean else 0)

    def correct(self):
        mm = np.empty(4)
        self.L.emul_correct(self.h, mm.ctypes.data_as(_dp))
        return mm

    def apply(self, op, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.zeros(self.n)
        self.L.emul_apply(self.h, op, x.ctypes.data_as(_dp), y.ctypes.data_as(_dp))
        return y

    def phi_ilu_w(self):
        """(w, (w, w)): w = ILU(0)(LHS_phi) * 1, what csrc/ns_hostgeom.h builds for outflow domains (zeros without an outflow edge)."""
        w = np.zeros(self.n)
        ww = self.L.emul_phi_ilu_w(self.h, w.ctypes.data_as(_dp))
        return w, ww

    def solve(self, which, rtol=1e-8, max_cycles=100):
        rel = C.c_double()
        n = self.L.emul_solve(self.h, which, rtol, max_cycles, C.byref(rel))
        return n, rel.value

    def poisson_sweeps(self, n):
        r = C.c_double()
        self.L.emul_poisson_sweeps(self.h, n, C.byref(r))
        return r.value

    def step(self, n=1, rtol=1e-8, max_cycles=100):
        mm = np.empty(4)
        cyc = (C.c_int * 3)()
        worst = self.L.emul_step(self.h, n, rtol, max_cycles, mm.ctypes.data_as(_dp), cyc)
        return mm, worst, list(cyc)


# ------------------------------------------------------------------------------------------------------------------
def golden_1024_gaps(u, v, phi):
    """Relative gaps of full 1024 x 1024 fields (compact-id order) to tests/golden/cavity1024_re1000_3steps.npz (written by
    tests/golden/make_golden_1024.py from the oracle at rtol 1e-12): sampled cells, six full lines, four functionals per field."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "cavity1024_re1000_3steps.npz"))
    n, stride, lines = 1024, int(z["stride"]), [int(k) for k in z["lines"]]
    wgt = np.cos(0.37 * np.arange(n * n) ** 0.5)
    out = {}
    for name, a in (("u", u), ("v", v), ("phi", phi - phi.mean())):
        f = a.reshape(n, n)
        scale = np.abs(z[f"{name}_cols"]).max()
        fun = np.array([a.sum(), (a * a).sum(), np.abs(a).max(), (a * wgt).sum()])
        ref = z[f"{name}_fun"]
        out[name] = dict(
            sub=float(np.abs(f[::stride, ::stride] - z[f"{name}_sub"]).max() / scale),
            cols=float(np.abs(f[lines, :] - z[f"{name}_cols"]).max() / scale),
            rows=float(np.abs(f[:, lines] - z[f"{name}_rows"]).max() / scale),
            sum=float(abs(fun[0] - ref[0]) / (n * n * scale)), sumsq=float(abs(fun[1] - ref[1]) / ref[1]),
            absmax=float(abs(fun[2] - ref[2]) / ref[2]), weighted=float(abs(fun[3] - ref[3]) / (n * n * scale)))
    return out, z["minmax"], float(z["dt"]), float(z["re"]), int(z["steps"])
