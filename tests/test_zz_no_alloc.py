"""nsgpu_minmax_noalloc (include/nsgpu.h): the entry point that allocates nothing on the device -- scratch in a file-scope static
__device__ array, address taken once per device with cudaGetSymbolAddress.  Collected last on purpose.
Checked against the existing entry point for the same quantity: nsgpu_correct's VecMin / VecMax x 4 (FS.cpp:554-557) with phi = 0, where
u = u*, v = v* -- exact equality (the tolerance the Tier-A min/max test uses), on two calls in a row that share the static scratch
(the second one shorter, so stale contents of the first would show)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_minmax_noalloc_equals_correct_entry_point_two_calls_in_a_row():
    import navierstokessolver_b200 as ns
    from navierstokessolver_b200 import capi
    rng = np.random.default_rng(2024)
    for ny, nx in ((300, 500), (96, 70)):              # second call: fewer cells than the first
        n = ny * nx
        gpu = ns.GpuFluidSolver(ns.GridDesc.cavity(ny, nx=nx, lx=nx / float(ny)), 0.01, 100.0)
        us, vs = rng.standard_normal(n), 3.0 * rng.standard_normal(n) + 0.5
        gpu.set(ns.USTAR, us)
        gpu.set(ns.VSTAR, vs)
        gpu.set(ns.PHI, np.zeros(n))
        info = gpu.correct()
        want = (info.umin, info.umax, info.vmin, info.vmax)
        assert want == (us.min(), us.max(), vs.min(), vs.max())
        got = capi.minmax_noalloc(us, vs)
        assert got == want
        gpu.close()
    # NaN propagation as in the time step, and the size limit
    us[7] = np.nan
    got = capi.minmax_noalloc(us, vs)
    assert got == (-np.inf, np.inf, -np.inf, np.inf)
    nmax = capi.lib.nsgpu_noalloc_max_cells()
    assert nmax == 1 << 20
    big = np.zeros(nmax + 1)
    with pytest.raises(ns.NsGpuError) as e:
        capi.minmax_noalloc(big, big)
    assert e.value.code == -1 and "exceeds the static scratch" in str(e.value)
    assert capi.minmax_noalloc(big[:nmax], big[:nmax]) == (0.0, 0.0, 0.0, 0.0)     # the maximum itself is accepted (-0.0 == 0.0)
