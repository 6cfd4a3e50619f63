"""Short program for ncu: `warm` warm-up steps then `k` time steps of the n x n cavity (Re 1000, dt = 0.25/n)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import navierstokessolver_b200 as ns
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
warm = int(sys.argv[2]) if len(sys.argv) > 2 else 2
k = int(sys.argv[3]) if len(sys.argv) > 3 else 1
s = ns.GpuFluidSolver(ns.GridDesc.cavity(n), 0.25 / n, 1000.0)
i0 = s.step(warm)
print("warm-up launches", i0.kernel_launches)
info = s.step(k)
print("steps", k, "ms/step", info.ms_total / k, "launches", info.kernel_launches, "it", info.it_u, info.it_v, info.it_phi)
