"""Ad-hoc GPU probe: fine-level sweep throughput and time-step cost (not the bench contract)."""
import sys, time, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import navierstokessolver_b200 as ns

for n in [int(a) for a in sys.argv[1:]] or [1024, 4096]:
    g = ns.GridDesc.cavity(n)
    s = ns.GpuFluidSolver(g, 0.25 / n, 1000.0)
    rng = np.random.default_rng(20240611)
    b = rng.uniform(-1, 1, g.n_cells); b -= b.mean()
    s.set(ns.RHS_PHI, b)
    s.poisson_sweeps(20, False)
    best = 1e9
    for _ in range(5):
        _, ms = s.poisson_sweeps(50, False)
        best = min(best, ms / 50)
    gc = n * n / best / 1e6
    print(json.dumps({"n": n, "sweep_ms": best, "gcell_updates_per_s": gc, "gbs_24B": gc * 24, "levels": s.num_levels}))
    s.set(ns.PHI, np.zeros(g.n_cells)); s.set(ns.RHS_PHI, np.zeros(g.n_cells))
    s.step(3)
    t = time.time(); info = s.step(10); dt = time.time() - t
    print(json.dumps({"n": n, "ms_per_step": info.ms_total / 10, "wall_ms_per_step": dt * 100, "it": [info.it_u, info.it_v, info.it_phi],
                      "res": [info.res_u, info.res_v, info.res_phi], "launches_per_step": info.kernel_launches / 10, "status": info.status,
                      "minmax": [info.umin, info.umax, info.vmin, info.vmax]}))
    s.close()
