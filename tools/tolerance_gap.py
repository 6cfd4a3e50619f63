"""Measures the field gap between the GPU path and the reference's own CPU solver (oracle/_ref) as a function of
the GPU's true-residual tolerance, next to the reference's own solver noise floor (its rtol 1e-8 run against its
rtol 1e-12 run).  This is the documented replacement for the "lexicographic -> red-black ordering gap" of
BASELINE.json: the reference has no SOR at all, its solves are PETSc Krylov.
TEST INFRASTRUCTURE: imports the oracle; never part of the product path."""
import json, os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import navierstokessolver_b200 as ns
import refapi
from refapi import CASES, U, V, PHI

def rel(a, b):
    return float(np.abs(a - b).max() / np.abs(b).max())

def run(grid, sim, steps, vel_solver):
    os.chdir(tempfile.mkdtemp())
    exact = refapi.RefSolver(f"{CASES}/{grid}", f"{CASES}/{sim}"); exact.set_rtol(1e-12); exact.step(steps)
    stock = refapi.RefSolver(f"{CASES}/{grid}", f"{CASES}/{sim}"); stock.step(steps)          # the reference as shipped: rtol 1e-8
    want = {k: exact.get(k) for k in (U, V, PHI)}
    out = {"case": grid, "steps": steps, "vel_solver": {0: "auto", 1: "multigrid", 2: "sor"}[vel_solver],
           "reference_rtol1e-8_vs_reference_rtol1e-12": {n: rel(stock.get(k), want[k]) for n, k in (("u", U), ("v", V), ("phi", PHI))}}
    for rtol in (1e-8, 1e-10, 1e-12):
        g = ns.GpuFluidSolver(ns.GridDesc(**exact.desc_kwargs()), exact.dt, exact.re, rtol_u=rtol, rtol_phi=rtol, vel_solver=vel_solver)
        info = g.step(steps)
        out["gpu_rtol%g_vs_reference_rtol1e-12" % rtol] = dict({n: rel(g.get(k), want[k]) for n, k in (("u", U), ("v", V), ("phi", PHI))},
                                                               status=info.status, it=[info.it_u, info.it_v, info.it_phi])
        g.close()
    return out

if __name__ == "__main__":
    res = [run("cavity_128.grid", "cavity_re1000_128.sim", 20, 0), run("cavity_128.grid", "cavity_re1000_128.sim", 20, 1),
           run("cavity_64.grid", "cavity_re100.sim", 50, 0), run("channel.grid", "channel.sim", 10, 0), run("step.grid", "step.sim", 10, 0)]
    print(json.dumps(res, indent=1))
