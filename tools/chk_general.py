import os, sys, tempfile, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import navierstokessolver_b200 as ns
import refapi
from refapi import CASES, U, V, PHI
os.chdir(tempfile.mkdtemp())
for grid, sim in (("channel.grid", "channel.sim"), ("step.grid", "step.sim")):
    ref = refapi.RefSolver(f"{CASES}/{grid}", f"{CASES}/{sim}"); ref.set_rtol(1e-12); ref.step(5)
    for solver in (0, 1):
        for vs in (0, 1):
            try:
                g = ns.GpuFluidSolver(ns.GridDesc(**ref.desc_kwargs()), ref.dt, ref.re, rtol_u=1e-10, rtol_phi=1e-10, solver=solver, vel_solver=vs)
                info = g.step(5)
                gaps = [float(np.abs(g.get(k) - ref.get(k)).max() / np.abs(ref.get(k)).max()) for k in (U, V, PHI)]
                print(grid, "solver", solver, "vel_solver", vs, "status", info.status, "it", [info.it_u, info.it_v, info.it_phi], "gaps", ["%.1e" % x for x in gaps], flush=True)
                g.close()
            except Exception as e:
                print(grid, solver, vs, "EXC", e, flush=True)
