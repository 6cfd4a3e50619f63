"""Short program for ncu: a few fine-level Poisson sweeps (k_smooth_rb) on an n x n cavity, nothing else."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import navierstokessolver_b200 as ns
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
k = int(sys.argv[2]) if len(sys.argv) > 2 else 6
g = ns.GridDesc.cavity(n)
s = ns.GpuFluidSolver(g, 0.25 / n, 1000.0)
b = np.random.default_rng(20240611).uniform(-1, 1, g.n_cells); b -= b.mean()
s.set(ns.RHS_PHI, b)
r, ms = s.poisson_sweeps(k, True)
print("sweeps", k, "ms/sweep", ms / k, "residual", r)
