#!/usr/bin/env python3
"""Experiment: initial guess of the phi solve.  The reference warm-starts from phi^n (FS.cpp:54); linear / quadratic extrapolation
in time from the previous solutions is compared here through the per-phase C ABI (the guess is set from the host, so only the
cycle counts mean something, not the times).   python tools/phi_guess_experiment.py [n] [steps] [skip]"""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import navierstokessolver_b200 as ns

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 14
skip = int(sys.argv[3]) if len(sys.argv) > 3 else 0
for order in (0, 1, 2):
    s = ns.GpuFluidSolver(ns.GridDesc.cavity(n), 0.25 / n, 1000.0)
    if skip:
        s.step(skip)
    hist, its, res0 = [s.get(ns.PHI)], [], []
    for k in range(steps):
        s.rhs_build(); s.solve_velocity(); s.div_rhs(True)
        if order == 1 and len(hist) >= 2:
            s.set(ns.PHI, 2.0 * hist[-1] - hist[-2])
        elif order == 2 and len(hist) >= 3:
            s.set(ns.PHI, 3.0 * hist[-1] - 3.0 * hist[-2] + hist[-3])
        info = s.solve_phi()
        its.append(int(info.it_phi))
        s.correct()
        hist.append(s.get(ns.PHI)); hist = hist[-3:]
    print(json.dumps({"n": n, "skip": skip, "extrapolation_order": order, "phi_cycles_per_step": its}), flush=True)
    s.close()
