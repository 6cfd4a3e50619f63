"""GPU parity tests added in round 2 ("What's weak" 1): the default tolerance, long runs into developed flow, the
random-domain sweep, the routes that had only run on the host emulation (stretched grids on b - c w, uncoarsenable grids), the
non-blocking export, checkpoints on the Krylov routes.  All through the C ABI (ctypes), oracle = the unmodified reference on the
PETSc stand-in (oracle/_ref/libnsref.so, prebuilt: /root/reference is not read at run time)."""
import os
import time

import numpy as np
import pytest

import refapi
from refapi import CASES, U, V, PHI, CONV_U0, CONV_V0, GRADPHI_X, GRADPHI_Y

pytestmark = pytest.mark.gpu


def _gpu(ref, **kw):
    import navierstokessolver_b200 as ns
    return ns.GpuFluidSolver(ns.GridDesc(**ref.desc_kwargs()), ref.dt, ref.re, **kw)


def _rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def test_default_tolerance_gap_is_the_documented_one(workdir):
    """What bench.py and bin/main run: true-residual rtol 1e-8 (the reference's KSP setting, FS.cpp:68,80).  DESIGN.md section 5 table:
    after 20 steps of cavity 128^2 Re 1000 the GPU fields are within 5e-8 of the reference converged to 1e-12 (measured 1e-10 ... 5e-10;
    the reference AS SHIPPED is at 2e-10 / 3e-10 / 5e-8 from its own converged answer)."""
    ref = refapi.RefSolver(f"{CASES}/cavity_128.grid", f"{CASES}/cavity_re1000_128.sim")
    ref.set_rtol(1e-12)
    gpu = _gpu(ref)                      # default tolerances
    ref.step(20)
    info = gpu.step(20)
    assert info.status == 0 and max(info.res_u, info.res_v, info.res_phi) <= 1e-8
    gaps = {k: _rel(gpu.get(k), ref.get(k)) for k in (U, V, PHI)}
    print("default-rtol gaps after 20 steps of cavity 128^2:", gaps)
    assert max(gaps.values()) <= 5e-8, gaps


@pytest.mark.parametrize("n,sim,steps", [(64, "cavity_re100.sim", 200), (64, "cavity_re1000_64.sim", 200), (128, "cavity_re1000_128.sim", 200),
                                         (256, "cavity_re1000_256.sim", 30)])
def test_long_runs_into_developed_flow(n, sim, steps, workdir):
    """SURVEY.md 8(d) parity runs: hundreds of steps (the primary vortex has formed; every earlier whole-step test stops after 8 steps from
    rest).  GPU at rtol 1e-10 and at the default 1e-8 against the oracle at 1e-12; the reference's observable -- the printed min/max table,
    FS.cpp:559-560 -- is compared at every 20th step.  Tolerances: 5e-8 at 1e-10 (the host emulation of the same loops measures 1e-8 on u
    at Re 100 after 200 steps), 5e-7 at 1e-8 (emulation: 6e-8)."""
    ref = refapi.RefSolver(f"{CASES}/cavity_{n}.grid", f"{CASES}/{sim}")
    ref.set_rtol(1e-12)
    tight, loose = _gpu(ref, rtol_u=1e-10, rtol_phi=1e-10), _gpu(ref)
    done = 0
    while done < steps:
        k = min(20, steps - done)
        mm = ref.step(k)
        for g, tol in ((tight, 5e-8), (loose, 5e-7)):
            info = g.step(k)
            assert info.status == 0, (done, info.status)
            assert np.allclose([info.umin, info.umax, info.vmin, info.vmax], mm, rtol=0, atol=tol), (done, tol)
        done += k
    for g, tol in ((tight, 5e-8), (loose, 5e-7)):
        gaps = {k: _rel(g.get(k), ref.get(k)) for k in (U, V, PHI)}
        print(n, sim, steps, "rtol", "1e-10" if g is tight else "1e-8", gaps)
        assert max(gaps.values()) <= tol, gaps


def test_developed_flow_restart_from_reference_state(workdir):
    """Developed flow as the INPUT: the oracle runs 150 steps, its complete state (u, v, phi, N_u, N_v, divPhi) is uploaded and both
    sides take 5 more steps -- the GPU's phases on a state it did not produce itself."""
    ref = refapi.RefSolver(f"{CASES}/cavity_64.grid", f"{CASES}/cavity_re1000_64.sim")
    ref.set_rtol(1e-12)
    ref.step(150)
    gpu = _gpu(ref, rtol_u=1e-10, rtol_phi=1e-10)
    for k in (U, V, PHI, CONV_U0, CONV_V0, GRADPHI_X, GRADPHI_Y):
        gpu.set(k, ref.get(k))
    for s in range(5):
        mm = ref.step(1)
        info = gpu.step(1)
        assert info.status == 0
        for k in (U, V, PHI):
            assert _rel(gpu.get(k), ref.get(k)) <= 2e-8, (s, k)
        assert np.allclose([info.umin, info.umax, info.vmin, info.vmax], mm, rtol=0, atol=1e-8)


def test_whole_steps_on_random_domains(workdir):
    """The 24 seeded random cases of tests/test_host_and_oracle.py (rectangles and L shapes, stretched segments, grids that cannot be
    coarsened, inlets / walls / one outflow edge) as whole time steps ON THE GPU: every solve converges, fields follow the reference
    (1e-6: the tolerance of the outflow tests, these cases mix all routes)."""
    from test_host_and_oracle import _random_case
    rng = np.random.default_rng(123)
    kinds, worst_gap = set(), 0.0
    for k in range(24):
        g, sim = _random_case(rng, workdir, k)
        ref = refapi.RefSolver(g, sim)
        ref.set_rtol(1e-12)
        gpu = _gpu(ref, rtol_u=1e-10, rtol_phi=1e-10, solver=0 if ref.n == len(ref.hx) * len(ref.hy) else 1)
        kinds.add((gpu.num_levels == 1, bool((ref.edge_type == 4).any()), ref.n == len(ref.hx) * len(ref.hy)))
        for s in range(2):
            ref.step(1)
            info = gpu.step(1)
            assert info.status == 0, (k, s, info.status, info.res_u, info.res_v, info.res_phi, info.it_phi)
            vel = max(np.abs(ref.get(U)).max(), np.abs(ref.get(V)).max())
            for q in (U, V):
                gap = np.abs(gpu.get(q) - ref.get(q)).max() / vel
                worst_gap = max(worst_gap, gap)
                assert gap <= 1e-6, (k, s, q, gap)
            gap = _rel(gpu.get(PHI), ref.get(PHI))
            worst_gap = max(worst_gap, gap)
            assert gap <= 1e-6, (k, s, gap)
        gpu.close()
    print("random domains: worst gap", worst_gap, "kinds", kinds)
    assert len(kinds) >= 4


def test_solver_free_phases_on_random_domains(workdir):
    """Tier A (1e-12) on the 30 seeded random domains / boundary-condition sets, GPU kernels against the reference's functions."""
    from test_host_and_oracle import _random_case
    import test_gpu_parity as tp
    rng = np.random.default_rng(77)
    for k in range(30):
        g, sim = _random_case(rng, workdir, k)
        ref = refapi.RefSolver(g, sim)
        gpu = _gpu(ref)
        r2 = np.random.default_rng(11 + k)
        for op in (0, 1):
            x = r2.standard_normal(ref.n)
            assert _rel(gpu.apply_operator(op, x), ref.csr(op) @ x) <= tp.TIER_A, (k, op)
        for q in (U, V, CONV_U0, CONV_V0, GRADPHI_X, GRADPHI_Y, PHI):
            a = 0.3 * r2.standard_normal(ref.n)
            ref.set(q, a)
            gpu.set(q, a)
        ref.construct_rhs_v()
        gpu.rhs_build()
        for q in (refapi.RHS_U, refapi.RHS_V, CONV_U0, CONV_V0):
            assert _rel(gpu.get(q), ref.get(q)) <= tp.TIER_A, (k, q)
        for q in (refapi.USTAR, refapi.VSTAR):
            a = r2.standard_normal(ref.n)
            ref.set(q, a)
            gpu.set(q, a)
        ref.construct_rhs_phi()
        ref.nullspace_remove()
        gpu.div_rhs(True)
        assert _rel(gpu.get(refapi.RHS_PHI), ref.get(refapi.RHS_PHI)) <= tp.TIER_A, k
        ref.correct()
        gpu.correct()
        for q in (U, V, GRADPHI_X, GRADPHI_Y):
            assert _rel(gpu.get(q), ref.get(q)) <= tp.TIER_A, (k, q)
        gpu.close()


@pytest.mark.parametrize("mode,ratio", [("1", 1.02), ("2", None), ("1", None)])
def test_stretched_cavity_multigrid_on_corrected_rhs(mode, ratio, workdir, monkeypatch):
    """NSGPU_RHS_CONSISTENT=1|2: the device-side multigrid (graphs, fused passes) on b - c w, c = (vol, b) / (vol, w) (k_vol_dot + one
    axpy) instead of the Krylov solve on the augmented operator -- same bordered system, same answer as the reference (Tier B).  Mode 1
    keeps multigrid only up to h_max / h_min = 2 (ratio 1.02 over 32 cells: 1.85); cases/cavity_stretched_64 (3.4) stays on the Krylov
    route there and takes multigrid in mode 2."""
    monkeypatch.setenv("NSGPU_RHS_CONSISTENT", mode)
    if ratio:
        (workdir / "s.grid").write_text("Vertices {\n0 0\n0 1\n1 1\n1 0\n}\nNx {\n0 0.5 32 %.10f\n0.5 1 32 %.10f\n}\nNy {\n0 0.5 32 %.10f\n0.5 1 32 %.10f\n}\n"
                                        % (ratio, 1 / ratio, ratio, 1 / ratio))
        ref = refapi.RefSolver(str(workdir / "s.grid"), f"{CASES}/cavity_stretched_64.sim")
        assert 1.5 < ref.hx.max() / ref.hx.min() <= 2.0
    else:
        ref = refapi.RefSolver(f"{CASES}/cavity_stretched_64.grid", f"{CASES}/cavity_stretched_64.sim")
    ref.set_rtol(1e-12)
    gpu = _gpu(ref, rtol_u=1e-10, rtol_phi=1e-10, max_cycles=400)
    expect_krylov = mode == "1" and ratio is None
    for s in range(4):
        mm = ref.step(1)
        info = gpu.step(1)
        assert info.status == 0, (s, info.status, info.res_phi, info.it_phi)
        assert bool(info.flags & 2) == expect_krylov
        for k in (U, V, PHI):
            assert _rel(gpu.get(k), ref.get(k)) <= 2e-8, (s, k, _rel(gpu.get(k), ref.get(k)))
        assert np.allclose([info.umin, info.umax, info.vmin, info.vmax], mm, rtol=0, atol=1e-8)


def test_uncoarsenable_grid_takes_the_krylov_route(workdir):
    """33 x 33 cells: one level only, the phi solve is BiCGStab preconditioned by the smoother (flags bit 1)."""
    (workdir / "c.grid").write_text("Vertices {\n0 0\n0 1\n1 1\n1 0\n}\nNx {\n0 1 33 -1\n}\nNy {\n0 1 33 -1\n}\n")
    ref = refapi.RefSolver(str(workdir / "c.grid"), f"{CASES}/cavity_re100.sim")
    ref.set_rtol(1e-12)
    gpu = _gpu(ref, rtol_u=1e-10, rtol_phi=1e-10)
    assert gpu.num_levels == 1
    for s in range(3):
        ref.step(1)
        info = gpu.step(1)
        assert info.status == 0 and (info.flags & 2), (s, info.status, info.flags)
        for q in (U, V, PHI):
            assert _rel(gpu.get(q), ref.get(q)) <= 2e-8, (s, q)


def test_strongly_stretched_cavity(workdir):
    """Wall refinement with ratio 1.1 (h_max / h_min = 19): plain multigrid diverges, the default (Krylov on the augmented operator) follows
    the reference."""
    r, inv = 1.1, 1.0 / 1.1
    (workdir / "s.grid").write_text("Vertices {\n0 0\n0 1\n1 1\n1 0\n}\nNx {\n0 0.5 32 %.10f\n0.5 1 32 %.10f\n}\nNy {\n0 0.5 32 %.10f\n0.5 1 32 %.10f\n}\n" % (r, inv, r, inv))
    ref = refapi.RefSolver(str(workdir / "s.grid"), f"{CASES}/cavity_stretched_64.sim")
    ref.set_rtol(1e-12)
    gpu = _gpu(ref, rtol_u=1e-10, rtol_phi=1e-10)
    for s in range(2):
        ref.step(1)
        info = gpu.step(1)
        assert info.status == 0, (s, info.status, info.res_phi, info.it_phi)
        for k in (U, V, PHI):
            assert _rel(gpu.get(k), ref.get(k)) <= 2e-8, (s, k)


@pytest.mark.parametrize("case", ["channel", "cavity_stretched_64"])
def test_checkpoint_restart_on_krylov_routes(case, workdir, tmp_path):
    """The warm-start shift of the augmented phi solve, dt and re are part of the checkpoint (header version 2): a restarted run continues
    bit for bit on outflow and stretched domains too; a file written with another dt is refused."""
    import navierstokessolver_b200 as ns
    ref = refapi.RefSolver(f"{CASES}/{case}.grid", f"{CASES}/{case}.sim")
    a = _gpu(ref)
    a.step(3)
    ck = tmp_path / "state.nsgpu"
    a.checkpoint_save(ck, step=3)
    ia = a.step(3)
    b = _gpu(ref)
    assert b.checkpoint_load(ck) == 3
    ib = b.step(3)
    assert (ia.it_u, ia.it_v, ia.it_phi) == (ib.it_u, ib.it_v, ib.it_phi)
    for k in (U, V, PHI, CONV_U0, CONV_V0, GRADPHI_X, GRADPHI_Y):
        assert np.array_equal(a.get(k), b.get(k)), k
    other = ns.GpuFluidSolver(ns.GridDesc(**ref.desc_kwargs()), ref.dt * 0.5, ref.re)
    with pytest.raises(ns.NsGpuError):
        other.checkpoint_load(ck)


@pytest.mark.parametrize("case", ["channel", "step", "cavity_stretched_64", "cavity_33"])
def test_krylov_solves_as_graphs_equal_direct_launches(case, workdir):
    """The Krylov route (outflow, masked, stretched, uncoarsenable grids) with its stopping test on the device: as one CUDA graph per
    (re)start (begin -> WHILE { BiCGStab iteration } -> true residual) and with the same kernels launched directly, the loop condition read
    back every iteration.  Same kernels in the same order: identical fields and iteration counts; the graph run must not have fallen back."""
    if case == "cavity_33":
        import navierstokessolver_b200 as ns
        mk = lambda **kw: ns.GpuFluidSolver(ns.GridDesc.cavity(33), 0.25 / 33, 100.0, **kw)
    else:
        ref = refapi.RefSolver(f"{CASES}/{case}.grid", f"{CASES}/{case}.sim")
        mk = lambda **kw: _gpu(ref, **kw)
    a, b = mk(use_graph=0), mk(use_graph=-1)
    t = {}
    for name, s in (("graph", a), ("direct", b)):
        s.step(2)
        t0 = time.perf_counter()
        info = s.step(4)
        t[name] = (time.perf_counter() - t0) / 4 * 1e3
        assert info.status == 0, (name, info.status)
        t[name + "_it"] = (info.it_u, info.it_v, info.it_phi)
        t[name + "_flags"] = info.flags
    print(f"\n{case}: ms/step graph {t['graph']:.2f}, direct {t['direct']:.2f}; iterations {t['graph_it']}; flags {t['graph_flags']}")
    assert t["graph_it"] == t["direct_it"]
    assert t["graph_flags"] & 2, "phi is expected on the Krylov route here"
    assert not (t["graph_flags"] & 1), "the graph run fell back to direct launches"
    for k in (U, V, PHI):
        assert np.array_equal(a.get(k), b.get(k)), k


def test_export_snapshot_matches_synchronous_downloads(workdir):
    """nsgpu_export_begin / _wait (pressure recovery on the device, FS.cpp:577-581): the snapshot equals the fields downloaded at the same
    step even though the time loop has moved on, Pr = phi - dt/(2Re) lapPhi exactly, lapPhi = LHS_phi phi (operator rows, Tier A), two
    snapshots may be in flight."""
    ref = refapi.RefSolver(f"{CASES}/step.grid", f"{CASES}/step.sim")
    for gpu, half in ((_gpu(ref, solver=1), ref.dt / (2 * ref.re)),):
        gpu.step(2)
        want = {k: gpu.get(k) for k in (U, V, PHI)}
        lap_ref = gpu.apply_operator(1, want[PHI])
        gpu.export_begin(2)
        gpu.step(2)                       # the loop goes on; the snapshot must not see it
        gpu.export_begin(4)
        want4 = {k: gpu.get(k) for k in (U, V, PHI)}
        gpu.step(1)
        tag, snap = gpu.export_wait()
        assert tag == 2
        for k, name in ((U, "u"), (V, "v"), (PHI, "phi")):
            assert np.array_equal(snap[name], want[k]), name
        assert np.array_equal(snap["lap"], lap_ref)
        assert np.array_equal(snap["pr"], want[PHI] - half * snap["lap"])
        A = ref.csr(1)
        assert _rel(snap["lap"], A @ want[PHI]) <= 1e-12
        tag, snap = gpu.export_wait()
        assert tag == 4
        for k, name in ((U, "u"), (V, "v"), (PHI, "phi")):
            assert np.array_equal(snap[name], want4[k]), name
        pr, lap = gpu.pressure()          # synchronous form = one snapshot waited for at once
        assert np.array_equal(pr, gpu.get(PHI) - half * lap)


@pytest.mark.parametrize("n", [1024, 2048])
def test_export_every_step_does_not_stall_the_loop(n, workdir):
    """`saveIter 1` must not stall the loop.  20 steps with a snapshot after every step (copies home overlapped,
    consumed one step late) against 20 plain steps.  Asserted: the snapshot kernel itself costs <= 5 % of a step, and the device time of
    the steps while snapshots travel home stays <= 1.5 x (measured 1.35-1.41 x: the per-step scalar readbacks -- three solve control
    blocks and the four min/max, written to mapped host memory -- share the device-to-host PCIe direction with 5 arrays per step of
    snapshot traffic; without the per-step readback, i.e. nsgpu_step(n > 1), nothing in the step touches PCIe).  The fields are the same
    with and without exports."""
    import ctypes as C
    import navierstokessolver_b200 as ns
    from navierstokessolver_b200 import capi
    steps = 20
    a = ns.GpuFluidSolver(ns.GridDesc.cavity(n), 0.25 / n, 1000.0)
    b = ns.GpuFluidSolver(ns.GridDesc.cavity(n), 0.25 / n, 1000.0)
    a.step(5); b.step(5)
    b.export_begin(0); b.export_wait()
    a.synchronize(); b.synchronize()
    t0 = time.perf_counter()
    dev_a = sum(a.step(1).ms_total for _ in range(steps))
    a.synchronize()
    wall_a = time.perf_counter() - t0
    snap_ms, dev_b, pending = 0.0, 0.0, 0
    ms = C.c_float(0.0)
    t0 = time.perf_counter()
    for s in range(steps):
        dev_b += b.step(1).ms_total
        if pending == 2:
            assert capi.lib.nsgpu_export_wait(b._h, None, None, None, None, None, None) == 0
            capi.lib.nsgpu_export_last_ms(b._h, C.byref(ms)); snap_ms += ms.value
            pending -= 1
        b.export_begin(s)
        pending += 1
    while pending:
        assert capi.lib.nsgpu_export_wait(b._h, None, None, None, None, None, None) == 0
        capi.lib.nsgpu_export_last_ms(b._h, C.byref(ms)); snap_ms += ms.value
        pending -= 1
    b.synchronize()
    wall_b = time.perf_counter() - t0
    print("export every step at %d^2: device ms/step %.3f -> %.3f (+ %.3f snapshot kernel) = %.2f x, wall %.3f -> %.3f ms/step"
          % (n, dev_a / steps, dev_b / steps, snap_ms / steps, (dev_b + snap_ms) / dev_a, 1e3 * wall_a / steps, 1e3 * wall_b / steps))
    assert snap_ms <= 0.05 * dev_a
    assert dev_b + snap_ms <= 1.5 * dev_a
    for k in (U, V, PHI):      # equal to rounding: an exported phi is normalised at once, a resident one lazily (nsgpu_ctx::lazy_mean)
        assert _rel(a.get(k), b.get(k)) <= 1e-12


def test_direct_launch_step_lists_the_solver_kernels(workdir):
    """use_graph = -1 (what __graft_entry__.smoke() also runs once): the same step without CUDA graphs, so that a profiler's launch list
    shows k_march* and k_coarse_tail; fields identical to the graph run."""
    import navierstokessolver_b200 as ns
    g = ns.GridDesc.cavity(128)
    a = ns.GpuFluidSolver(g, 0.25 / 128, 1000.0, use_graph=0)
    b = ns.GpuFluidSolver(g, 0.25 / 128, 1000.0, use_graph=-1)
    ia, ib = a.step(2), b.step(2)
    assert not (ia.flags & 1) and ia.kernel_launches == ib.kernel_launches
    for k in (U, V, PHI):
        assert np.array_equal(a.get(k), b.get(k))


@pytest.mark.parametrize("ny,nx", [(2048, 2048), (1000, 1030), (4096, 1024)])
def test_projection_kernels_random_fields_repeatable(ny, nx, workdir, monkeypatch):
    """k_div_march / k_correct_march read their inputs through a per-warp shared-memory ring refilled by bulk copies (cp.async.bulk +
    mbarrier).  A refill issued too early overwrites a slot some lane has not read yet -- a race that smooth flow fields hide (the
    neighbouring column is almost the same) and white noise does not: random u*, v*, phi on grids with hundreds of refills per warp,
    eight repeats each bit-identical, and equal to the one-thread-per-row kernels (NSGPU_PHYS_MARCH=0) to rounding."""
    import navierstokessolver_b200 as ns
    rng = np.random.default_rng(ny + nx)
    g = ns.GridDesc.cavity(ny, nx=nx, lx=nx / float(ny))
    fields = {k: rng.standard_normal(ny * nx) for k in (refapi.USTAR, refapi.VSTAR, PHI)}
    out = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("NSGPU_PHYS_MARCH", mode)
        gpu = ns.GpuFluidSolver(g, 0.25 / ny, 1000.0)
        runs = []
        for rep in range(8 if mode == "1" else 1):
            for k, a in fields.items():
                gpu.set(k, a)
            gpu.div_rhs(True)
            b = gpu.get(refapi.RHS_PHI)
            info = gpu.correct()
            runs.append((b, gpu.get(U), gpu.get(V), gpu.get(GRADPHI_X), gpu.get(GRADPHI_Y), (info.umin, info.umax, info.vmin, info.vmax)))
        for r in runs[1:]:
            for a, b in zip(r[:5], runs[0][:5]):
                assert np.array_equal(a, b)
            assert r[5] == runs[0][5]
        out[mode] = runs[0]
        gpu.close()
    for a, b in zip(out["1"][:5], out["0"][:5]):
        assert np.abs(a - b).max() <= 1e-13 * np.abs(b).max()
    assert out["1"][5] == out["0"][5]


@pytest.mark.parametrize("ny,nx", [(256, 256), (200, 130), (1000, 1024), (2048, 2048)])
def test_ring_march_equals_register_march(ny, nx, workdir, monkeypatch):
    """ns_march_ring.cuh (columns of x and b through a per-warp shared-memory ring filled by bulk copies) against ns_march.cuh (register
    ring): same arithmetic in the same order -- identical fields and iteration counts after whole time steps, on even, odd and ragged sizes
    (strips that end inside a warp, warps wholly past the last row)."""
    import navierstokessolver_b200 as ns
    out = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("NSGPU_MARCH_RING", mode)
        s = ns.GpuFluidSolver(ns.GridDesc.cavity(ny, nx=nx, lx=nx / float(ny)), 0.25 / ny, 1000.0)
        info = s.step(4)
        assert info.status == 0
        out[mode] = ((info.it_u, info.it_v, info.it_phi), {k: s.get(k) for k in (U, V, PHI)})
        s.close()
    assert out["0"][0] == out["1"][0]
    for k in (U, V, PHI):
        assert np.array_equal(out["0"][1][k], out["1"][1][k]), k


def test_cluster_tail_of_16_ctas(workdir, monkeypatch):
    """NSGPU_CTAIL_C=16 + NSGPU_CTAIL_CELLS=65536: the coarse tail starts at 256 x 256 on a (non-portable) cluster of 16 CTAs instead of at
    128 x 128 on 8.  Same algorithm, another order of the same sums nowhere: the levels a cluster holds are smoothed in shared memory instead of
    by marching passes -- the fields agree to rounding and the iteration counts are equal.  Skips when the GPU cannot schedule 16-CTA clusters."""
    import navierstokessolver_b200 as ns
    out = {}
    for mode in ("8", "16"):
        if mode == "16":
            monkeypatch.setenv("NSGPU_CTAIL_C", "16")
            monkeypatch.setenv("NSGPU_CTAIL_CELLS", "65536")
        s = ns.GpuFluidSolver(ns.GridDesc.cavity(1024), 0.25 / 1024, 1000.0)
        info = s.step(4)
        assert info.status == 0
        out[mode] = ((info.it_u, info.it_v, info.it_phi), {k: s.get(k) for k in (U, V, PHI)}, info.ms_total / 4)
        s.close()
    print("\nms/step: 8-CTA tail from 128^2", round(out["8"][2], 3), " 16-CTA tail from 256^2", round(out["16"][2], 3))
    assert out["8"][0] == out["16"][0]
    for k in (U, V, PHI):
        assert _rel(out["16"][1][k], out["8"][1][k]) <= 1e-9, k
