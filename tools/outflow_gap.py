#!/usr/bin/env python3
"""Whole time steps on the outflow (bcType 4) domains, GPU against the oracle (reference sources + PETSc stand-in at rtol 1e-12):
relative max-norm gap of u, v, phi per step, BiCGStab iterations of the augmented phi solve, time per step.  One JSON line per case.
  python tools/outflow_gap.py [steps]"""
import json, os, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import refapi
from refapi import CASES, U, V, PHI
import navierstokessolver_b200 as ns

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 8
os.chdir(tempfile.mkdtemp())
for case in ("channel", "step"):
    for rtol in (1e-8, 1e-10):
        ref = refapi.RefSolver(f"{CASES}/{case}.grid", f"{CASES}/{case}.sim")
        ref.set_rtol(1e-12)
        gpu = ns.GpuFluidSolver(ns.GridDesc(**ref.desc_kwargs()), ref.dt, ref.re, rtol_u=rtol, rtol_phi=rtol)
        rows = []
        for s in range(steps):
            ref.step(1)
            t0 = time.perf_counter()
            info = gpu.step(1)
            ms = 1e3 * (time.perf_counter() - t0)
            gap = [float(np.abs(gpu.get(k) - ref.get(k)).max() / np.abs(ref.get(k)).max()) for k in (U, V, PHI)]
            rows.append({"step": s + 1, "status": int(info.status), "it": [info.it_u, info.it_v, info.it_phi],
                         "res": [info.res_u, info.res_v, info.res_phi], "gap_u_v_phi": gap, "wall_ms": ms})
        print(json.dumps({"case": case, "cells": int(ref.n), "gpu_rtol": rtol, "reference_rtol": 1e-12, "steps": rows}), flush=True)
        gpu.close()
