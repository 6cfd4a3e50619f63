"""Per-pass timing table of the time step (NSGPU_TRACE=1, direct launches): python tools/trace_step.py [n] [steps]
key = level * 1000 + nsw * 100 + res * 10 + prolong * 2 + push (+50000: SOR passes of the velocity solves); 99999 = coarse tail."""
import os, sys
os.environ["NSGPU_TRACE"] = "1"
os.environ["NSGPU_GRAPH"] = "0"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import navierstokessolver_b200 as ns
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
k = int(sys.argv[2]) if len(sys.argv) > 2 else 5
s = ns.GpuFluidSolver(ns.GridDesc.cavity(n), 0.25 / n, 1000.0)
s.step(3)
info = s.step(k)
print("steps", k, "ms/step", info.ms_total / k, "it", info.it_u, info.it_v, info.it_phi, "phases", s.phase_ms(), file=sys.stderr)
s.close()
