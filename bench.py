#!/usr/bin/env python3
"""bench.py -- the time-step loop of the lid-driven cavity on N B200s (contract: see DESIGN.md "Measurement").

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torch.distributed.run, one rank per GPU)
  python bench.py --impl reference --steps K --warmup W    (the reference's own CPU code on the host cores)

A "step" is one pass of FluidSolver::Solve()'s loop body (reference FS.cpp:546-557): RHS build, two Helmholtz
solves, divergence + mean removal, Poisson solve, projection, min/max -- every solve converged to a relative true
residual of 1e-8 (the reference's KSP tolerance).  Workload at N = 1: lid-driven cavity, Re = 1000, 4096 x 4096 cells
(BASELINE.json configs[2]); weak scaling: every GPU owns a 4096-column slab, the domain is (4096 N) x 4096.

  value  = cells x time steps / second, state resident in HBM, device time (CUDA events on the solver's stream,
           max over ranks), in Gcell-updates/s (one cell-update = one cell advanced one time step);
           "timesteps_per_s" and "poisson_gcell_updates_per_s" (the fine-level red-black sweep alone) ride along.
  e2e    = same metric through the C ABI with HOST buffers (nsgpu_step_host): every step uploads the complete state of the
           time loop from pinned host memory -- u, v, phi and the two convective terms; divPhi is GradP(phi) and is rebuilt on
           the device instead of being shipped -- runs one time step and downloads the five arrays again; the copies overlap
           the phases.  (--e2e-arrays 7 ships divPhi as well, like a caller that mirrors every reference Vec.)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# The JSON line is the only thing this program writes to its stdout: file descriptor 1 is pointed at stderr for everything else
# (NCCL prints its version banner on stdout when NCCL_DEBUG is set), the line itself goes to a duplicate of the original stdout.
_JSON_OUT = None


def _isolate_stdout():
    global _JSON_OUT
    if _JSON_OUT is None:
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _JSON_OUT if _JSON_OUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


METRIC = "timestep_gcell_updates_per_s"
UNIT = "Gcell-updates/s"
SWEEP_BYTES_PER_UPDATE = 24.0     # read phi, read b, write phi


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md, 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) > 2 + k and r[2 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons, "samples": len(self.rows)}


def reference_arm(args, rank):
    """The reference's own CPU implementation (unmodified sources + PETSc stand-in, oracle/_ref) on a bounded sample."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import tempfile
    import refapi
    n = args.ref_n
    cases = os.path.join(ROOT, "cases")
    os.chdir(tempfile.mkdtemp())
    ref = refapi.RefSolver(f"{cases}/cavity_{n}.grid", f"{cases}/cavity_re1000_{n}.sim")
    ref.step(args.warmup)
    t0 = time.perf_counter()
    ref.step(args.steps)
    dt = time.perf_counter() - t0
    value = n * n * args.steps / dt / 1e9
    ph = ref.phase_seconds()
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "lid-driven cavity Re=1000 from rest, dt=0.25/n, bounded sample %dx%d of the 4096x4096 workload" % (n, n),
                       "solver": "reference FluidSolver (unmodified) on the PETSc-subset stand-in: ILU(0)-GMRES(30) + ILU(0)-BiCGStab(2), rtol 1e-8"},
            "timesteps_per_s": args.steps / dt,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "reference",
                             "sample": "%d steps of cavity %dx%d after %d warm-up steps; the reference is strictly serial (PETSC_COMM_SELF)" % (args.steps, n, n, args.warmup)},
            "phase_seconds": dict(zip(["rhs", "helmholtz", "div", "poisson", "correct"], [float(x) for x in ph])),
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def cpu_baseline_sample(n=512, steps=2):
    """Oracle (reference sources + PETSc stand-in) on the host cores, bounded sample; reported, not the target."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import tempfile
    import refapi
    cases = os.path.join(ROOT, "cases")
    cwd = os.getcwd()
    os.chdir(tempfile.mkdtemp())
    try:
        ref = refapi.RefSolver(f"{cases}/cavity_{n}.grid", f"{cases}/cavity_re1000_{n}.sim")
        t0 = time.perf_counter()
        ref.step(steps)
        dt = time.perf_counter() - t0
        its = ref.ksp_stats(1)["iterations_total"]
    finally:
        os.chdir(cwd)
    return {"value": n * n * steps / dt / 1e9, "unit": UNIT, "cores": 1, "kind": "reference",
            "sample": "first %d steps of cavity %dx%d Re=1000 from rest (%.1f s, %d BiCGStab(2) iterations); host has %d cores, the reference uses 1"
                      % (steps, n, n, dt, its, os.cpu_count()),
            "timesteps_per_s": steps / dt}


def timed_case(ns, torch, dist, rank, world, local_rank, nccl_id, ny, nxl, args, steps, warmup, e2e=False, fixed_cycles=0):
    """One more cavity (Re 1000 from rest, dt = 0.25/ny, ny rows, nxl columns per rank) timed like the headline: `warmup` untimed
    steps, `steps` timed ones, device time, max over ranks.  fixed_cycles > 0: every solve runs exactly that many iterations."""
    import numpy as np
    gnx = nxl * world
    if world > 1:   # every context has its own communicator: a fresh NCCL id, agreed on by all ranks
        from navierstokessolver_b200 import capi
        box = [capi.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        nccl_id = box[0]
    try:
        grid = ns.GridDesc.cavity(ny, nx=gnx, lx=gnx / float(ny))
        kw = dict(device=local_rank, rank=rank, world=world, nccl_id=nccl_id, solver=args.solver, use_graph=0 if args.graph else -1)
        if fixed_cycles:
            kw.update(rtol_u=1e-300, rtol_phi=1e-300, max_cycles=fixed_cycles)
        else:
            kw.update(rtol_u=args.rtol, rtol_phi=args.rtol)
        s = ns.GpuFluidSolver(grid, 0.25 / ny, args.re, **kw)
    except ns.NsGpuError as e:
        return {"ny": ny, "nx": gnx, "unavailable": str(e)}

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def mx(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    s.step(warmup)
    sync_all()
    info = s.step(steps)
    ms = mx(info.ms_total) / steps
    rec = {"ny": ny, "nx": gnx, "columns_per_gpu": nxl, "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms,
           "gcell_updates_per_s": gnx * ny / (ms * 1e-3) / 1e9, "timesteps_per_s": 1e3 / ms,
           "iterations_last_step": {"u": info.it_u, "v": info.it_v, "phi": info.it_phi},
           "phase_ms_last_step": dict(zip(["rhs_build", "velocity_solves", "div_rhs", "phi_solve", "correct"], s.phase_ms())),
           "status": int(info.status), "levels": s.num_levels}
    if fixed_cycles:
        rec["fixed_iterations_per_solve"] = fixed_cycles
        rec["fine_sweeps_per_step"] = {"phi": 4 * fixed_cycles, "u": 4 * fixed_cycles, "v": 4 * fixed_cycles}
        rec["note"] = "status -5 (NOT_CONVERGED) is expected: the tolerance is unreachable on purpose"
    if e2e:
        keys = [ns.U, ns.V, ns.PHI, ns.CONV_U0, ns.CONV_V0]
        host = {k: torch.empty(s.n_local, dtype=torch.float64).pin_memory().numpy() for k in keys}
        for k in keys:
            host[k][:] = s.get(k)
        s.step_host(host)
        sync_all()
        t0 = time.perf_counter()
        for _ in range(steps):
            s.step_host(host)
        s.synchronize()
        rec["e2e_ms_per_step"] = 1e3 * mx(time.perf_counter() - t0) / steps
        rec["e2e_gcell_updates_per_s"] = gnx * ny / (rec["e2e_ms_per_step"] * 1e-3) / 1e9
    s.close()
    torch.cuda.empty_cache()
    return rec


def parity_vs_one_rank(ns, capi, torch, dist, rank, world, local_rank, nccl_id, args):
    """N x-slabs against ONE rank on the same problems (4 steps, bench tolerances, graphs as in the timed run): the driver's test box
    has one GPU, so this is where N-rank correctness is checked under the driver's clock.  Every rank needs its own NCCL id per
    context: the id of the timed context is reused (a communicator per context, created and destroyed in the same order on all ranks)."""
    import numpy as np
    out = {"cases": [], "steps": 4}
    worst = {"u": 0.0, "v": 0.0, "phi": 0.0}
    its_equal = True
    for ny, nxl in ((256, 128), (130, 66), (512, 256)):
        gnx = nxl * world
        grid = ns.GridDesc.cavity(ny, nx=gnx, lx=gnx / float(ny))
        box = [capi.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        kw = dict(rtol_u=args.rtol, rtol_phi=args.rtol, use_graph=0 if args.graph else -1)
        s, err = None, ""
        try:
            s = ns.GpuFluidSolver(grid, 0.25 / ny, args.re, device=local_rank, rank=rank, world=world, nccl_id=box[0], **kw)
        except ns.NsGpuError as e:   # deterministic (plan / arguments): every rank sees it; the flag below keeps the collectives in step anyway
            err = str(e)
        okt = torch.tensor([1 if s is not None else 0], dtype=torch.int32, device="cuda")
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        if int(okt.item()) == 0:
            if s is not None:
                s.close()
            if rank == 0:
                out["cases"].append({"grid": [gnx, ny], "unavailable": err or "context creation failed on another rank"})
            continue
        info = s.step(4)
        mine = {k: torch.from_numpy(s.get(k)).cuda() for k in (ns.U, ns.V, ns.PHI)}
        full = {}
        for k, t in mine.items():
            parts = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
            dist.gather(t, parts, dst=0)
            if rank == 0:
                full[k] = torch.cat(parts).cpu().numpy()
        flags = int(info.flags)
        s.close()
        if rank == 0:
            one = ns.GpuFluidSolver(grid, 0.25 / ny, args.re, device=local_rank, **kw)
            i1 = one.step(4)
            gaps = {}
            for k, name in ((ns.U, "u"), (ns.V, "v"), (ns.PHI, "phi")):
                ref = one.get(k)
                gaps[name] = float(np.abs(full[k] - ref).max() / max(np.abs(ref).max(), 1e-300))
                worst[name] = max(worst[name], gaps[name])
            same = [info.it_u, info.it_v, info.it_phi] == [i1.it_u, i1.it_v, i1.it_phi]
            its_equal = its_equal and same
            out["cases"].append({"grid": [gnx, ny], "gaps": gaps, "iterations": [info.it_u, info.it_v, info.it_phi],
                                 "iterations_1gpu": [i1.it_u, i1.it_v, i1.it_phi], "status": [int(info.status), int(i1.status)],
                                 "p2p": bool(flags & 4), "graph_fallback": bool(flags & 1)})
            one.close()
        dist.barrier()
    out.update(worst)
    out["iterations_equal"] = its_equal
    out["tolerance"] = "relative max-norm gap of u, v, phi after 4 steps, rtol %g on both sides; summation order is the only difference" % args.rtol
    return out


def main_on_n_gpus(n):
    """The drop-in driver on n GPUs of this box: `bin/main <grid> <sim>` with NSGPU_GPUS=n against the same command on one GPU --
    stdout and FlowData CSVs compared (tests/test_gpu_parity.py holds the comparison; the driver's test box has one GPU)."""
    import pathlib
    import tempfile
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    try:
        import test_gpu_parity as tp
        wd = pathlib.Path(tempfile.mkdtemp())
        t0 = time.perf_counter()
        one = tp.run_main_on_gpus(1, wd, steps=20)
        t1 = time.perf_counter()
        many = tp.run_main_on_gpus(n, wd, steps=20)
        t2 = time.perf_counter()
        worst = tp.compare_main_runs(one, many)
        return {"gpus": n, "case": "cavity 128x128, 20 steps, exports every 5", "worst_difference": worst,
                "ok": bool(worst <= 2e-6), "wall_s": {"1": t1 - t0, str(n): t2 - t1}}
    except Exception as e:   # reported, never fatal for the bench line
        return {"gpus": n, "ok": False, "error": repr(e)[:400]}

def ours(args, rank, world, local_rank):
    import numpy as np
    import navierstokessolver_b200 as ns
    from navierstokessolver_b200 import capi
    import torch
    import torch.distributed as dist

    torch.cuda.set_device(local_rank)
    nccl_id = None
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
        box = [capi.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        nccl_id = box[0]

    ny, nxl = args.ny, args.nx_per_gpu
    gnx = nxl * world
    n = ny                                     # dt follows the cell size like the reference inputs: 0.25 * h
    grid = ns.GridDesc.cavity(ny, nx=gnx, lx=gnx / float(ny))
    s = ns.GpuFluidSolver(grid, 0.25 / n, args.re, device=local_rank, rank=rank, world=world, nccl_id=nccl_id,
                          rtol_u=args.rtol, rtol_phi=args.rtol, solver=args.solver, use_graph=0 if args.graph else -1)
    cells_local, cells_global = nxl * ny, gnx * ny

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- the time-step loop, state resident
    s.step(args.warmup)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    t0 = time.perf_counter()
    info = s.step(args.steps)
    s.synchronize()
    wall = time.perf_counter() - t0
    barrier()
    clocks = sampler.summary()
    ms = max_over_ranks(info.ms_total)
    wall = max_over_ranks(wall)
    value = cells_global * args.steps / (ms * 1e-3) / 1e9
    launches = int(info.kernel_launches)
    phase_ms = s.phase_ms()
    fine_sweeps = float(info.fine_sweeps)

    # ---- the dominant kernel alone: fine-level red-black Poisson sweep (k_march<1,false,0>), CUDA events on its stream
    s.poisson_sweeps(20, False)
    best = None
    for _ in range(5):
        barrier()
        _, sms = s.poisson_sweeps(args.sweeps, False)
        sms = max_over_ranks(sms) / args.sweeps
        best = sms if best is None else min(best, sms)
    sweep_gcups = cells_global / (best * 1e-3) / 1e9
    peak, peak_src = measured_peak_gbs()
    achieved = cells_local * SWEEP_BYTES_PER_UPDATE / (best * 1e-3) / 1e9
    # dram bytes per launch from the committed ncu capture -- only when the timed launch IS the captured configuration
    # (one rank, 4096 x 4096, non-pushing kernel variant); any other grid or rank count reports null
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "sweep_traffic.json")
    if world == 1 and ny == 4096 and nxl == 4096 and os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get("dram_bytes_per_launch")
        except Exception:
            traffic = None

    # ---- the passes the time step really runs on the fine level, each timed alone the same way (nsgpu_time_passes)
    inloop = {}
    if s.num_levels >= 2:
        names = {1: "k_march<2,false,1,*> first pass of a phi V-cycle: 2 fused red-black sweeps + residual restricted to level 1",
                 2: "k_march<2,true,2,*> last pass of a phi V-cycle: coarse correction added, 2 fused sweeps, |r|^2 reduced",
                 3: "k_march<2,false,0,*,SOR> red-black SOR double sweep of LHS_V (velocity solves)"}
        # algorithmic bytes per fine cell and pass: read x, read b, write x = 24; the restriction writes / the prolongation reads
        # one coarse double per 4 fine cells (+2)
        bpc = {1: 26.0, 2: 26.0, 3: 24.0}
        try:
            for kind in (1, 2, 3):
                s.time_passes(kind, 10)
                b_ = None
                for _ in range(3):
                    barrier()
                    t_ = max_over_ranks(s.time_passes(kind, 50)) / 50
                    b_ = t_ if b_ is None else min(b_, t_)
                ach = cells_local * bpc[kind] / (b_ * 1e-3) / 1e9
                inloop[str(kind)] = {"kernel": names[kind], "us_per_launch": b_ * 1e3, "bytes_per_cell": bpc[kind], "cell_updates_per_launch": 2 * cells_local,
                                     "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                                     "gcell_updates_per_s": 2 * cells_global / (b_ * 1e-3) / 1e9}
        except ns.NsGpuError as e:
            inloop = {"unavailable": str(e)}
    # ncu DRAM bytes of the in-loop passes, same rule as `traffic`
    ipath = os.path.join(ROOT, "profiles", "inloop_traffic.json")
    if world == 1 and ny == 4096 and nxl == 4096 and os.path.exists(ipath) and "unavailable" not in inloop:
        try:
            for k_, v_ in json.load(open(ipath)).items():
                if k_ in inloop:
                    inloop[k_]["traffic"] = v_.get("dram_bytes_per_launch")
        except Exception:
            pass

    # ---- end to end through the C ABI with host buffers (pinned), full state up, fields down, every step
    fields_up = [ns.U, ns.V, ns.PHI, ns.CONV_U0, ns.CONV_V0] + ([ns.GRADPHI_X, ns.GRADPHI_Y] if args.e2e_arrays == 7 else [])
    host = {k: torch.empty(s.n_local, dtype=torch.float64).pin_memory().numpy() for k in fields_up}
    for k in fields_up:
        host[k][:] = s.get(k)
    e2e_steps = max(2, min(args.steps, 5))
    s.step_host(host)   # one untimed step: copy streams, events and the solve graphs of this form are made on first use
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        s.step_host(host)   # the state arrays up, one time step, the same arrays down; copies overlap the phases (nsgpu_step_host)
    s.synchronize()
    e2e_wall = max_over_ranks(time.perf_counter() - t0)
    e2e_value = cells_global * e2e_steps / e2e_wall / 1e9
    h2d = 8 * s.n_local * len(fields_up)
    d2h = 8 * s.n_local * len(fields_up)

    n_levels = s.num_levels
    s.close()
    del host
    torch.cuda.empty_cache()

    extras = {}
    if not args.no_extras:
        # ---- (a) several ranks: the same small problems on N slabs and on ONE rank (rank 0's GPU), fields compared
        if world > 1:
            extras["parity_vs_1gpu"] = parity_vs_one_rank(ns, capi, torch, dist, rank, world, local_rank, nccl_id, args)
        # ---- (b) north-star configuration: 16384 x 16384, strong scaling (16384 / N columns per GPU), one code state
        extras["strong_16384"] = timed_case(ns, torch, dist, rank, world, local_rank, nccl_id, 16384, 16384 // world, args, steps=5, warmup=3)
        if world == 1:
            # ---- (c) the reference arm's own inputs (like for like), and the other size SURVEY 8(d) asks a time-step line for
            extras["same_config"] = [timed_case(ns, torch, dist, rank, world, local_rank, nccl_id, n_, n_, args, steps=args.steps, warmup=args.warmup, e2e=True)
                                     for n_ in (256, 512)]
            extras["size_1024"] = timed_case(ns, torch, dist, rank, world, local_rank, nccl_id, 1024, 1024, args, steps=args.steps, warmup=args.warmup)
            # ---- (d) fixed work: exactly 4 phi V-cycles and 4 x 4 SOR sweeps per velocity solve every step, whatever the residuals
            extras["fixed_work_4096"] = timed_case(ns, torch, dist, rank, world, local_rank, nccl_id, ny, nxl, args, steps=args.steps, warmup=3, fixed_cycles=4)

        # ---- (e) several GPUs behind the unchanged command line: bin/main with NSGPU_GPUS=N against NSGPU_GPUS=1 (rank 0 runs both)
        if world > 1:
            if rank == 0:
                extras["main_multi_gpu"] = main_on_n_gpus(world)
            dist.barrier()

    cpu = cpu_baseline_sample() if (rank == 0 and world == 1 and not args.no_cpu_baseline) else None

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": {"workload": "lid-driven cavity Re=%g from rest, %dx%d cells (%d columns per GPU), dt=0.25/%d" % (args.re, gnx, ny, nxl, n),
                           "parallelism": ("x-slabs, %d rank(s); solves: halo columns, norms and the replicated coarse rhs through NVLink peer memory from inside the kernels (NCCL only outside the solves)" % world) if world > 1 else "1 rank (no exchange)",
                           "solver": ("BiCGStab + V(2,2)" if args.solver else "phi: multigrid V(2,2), fused red-black GS passes; u*, v*: red-black SOR or multigrid (auto)")
                                     + ", true-residual rtol %g all solves, %s" % (args.rtol, "one CUDA graph per solve (device-side WHILE)" if args.graph else "direct launches"),
                           "levels": n_levels, "cache": "inputs larger than L2 (13 resident fields x %.0f MB per GPU)" % (cells_local * 8 / 1e6)},
                "timesteps_per_s": args.steps / (ms * 1e-3),
                "poisson_gcell_updates_per_s": sweep_gcups,
                "solver_iterations_last_step": {"u": info.it_u, "v": info.it_v, "phi": info.it_phi},
                "residuals_last_step": {"u": info.res_u, "v": info.res_v, "phi": info.res_phi},
                "fine_poisson_sweeps_per_step": fine_sweeps / args.steps,
                "host_wall_ms_per_step": 1e3 * wall / args.steps,
                "phase_ms_last_step": dict(zip(["rhs_build", "velocity_solves", "div_rhs", "phi_solve", "correct"], phase_ms)),
                "gpu_launches": launches,
                "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                        "ms_per_step": 1e3 * e2e_wall / e2e_steps,
                        "arrays": "%d state arrays each way per step (u, v, phi, N_u, N_v%s), pinned host memory, nsgpu_step_host"
                                  % (len(fields_up), ", divPhi x/y" if args.e2e_arrays == 7 else "; divPhi = GradP(phi) rebuilt on the device")},
                "roofline": {"bound": "hbm", "kernel": "k_march<1,false,0> (fine-level fused red-black Poisson sweep, register-marching)", "achieved": achieved, "peak": peak,
                             "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                             "bytes_per_cell_update": SWEEP_BYTES_PER_UPDATE, "us_per_launch": best * 1e3,
                             "cells_per_launch": cells_local, "timing": "CUDA events on the solver stream around %d back-to-back launches, best of 5" % args.sweeps,
                             "rhs": "physically shaped: phi and RHS_phi are the state after %d cavity time steps" % (args.warmup + args.steps)},
                "status": int(info.status), "flags": int(info.flags)}
        line["roofline_inloop"] = inloop
        line.update(extras)
        if cpu:
            line["cpu_baseline"] = cpu
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ny", type=int, default=4096)
    ap.add_argument("--nx-per-gpu", type=int, default=4096)
    ap.add_argument("--re", type=float, default=1000.0)
    ap.add_argument("--rtol", type=float, default=1e-8)
    ap.add_argument("--solver", type=int, default=0)
    ap.add_argument("--graph", type=int, default=1)
    ap.add_argument("--sweeps", type=int, default=100)
    ap.add_argument("--ref-n", type=int, default=256)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-arrays", type=int, default=5, choices=[5, 7])
    ap.add_argument("--no-extras", action="store_true", help="skip the sub-records (parity_vs_1gpu, strong_16384, same_config, size_1024, fixed_work_4096)")
    args = ap.parse_args()
    _isolate_stdout()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            sys.exit("bench.py --gpus %d must be launched with torch.distributed.run --nproc-per-node %d" % (args.gpus, args.gpus))
    ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
