"""How many V-cycles the phi solve needs as a function of its tolerance (cavity Re 1000, n x n, steps from rest): the convergence curve of the
cycle seen through the stopping test.   python tools/phi_cycle_curve.py [n] [steps]"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import navierstokessolver_b200 as ns
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 13
for rtol in (1e-5, 1e-6, 1e-7, 3e-8, 2e-8, 1e-8, 5e-9, 2e-9, 1e-9, 1e-10):
    s = ns.GpuFluidSolver(ns.GridDesc.cavity(n), 0.25 / n, 1000.0, rtol_phi=rtol)
    its, res = [], []
    for k in range(steps):
        info = s.step(1)
        its.append(info.it_phi); res.append(float("%.2e" % info.res_phi))
    print(json.dumps({"rtol_phi": rtol, "it_phi_per_step": its, "res_phi_per_step": res, "ms_last_step": round(info.ms_total, 3)}), flush=True)
    s.close()
