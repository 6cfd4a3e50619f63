#!/bin/bash
# compute-sanitizer on small grids, graphs off: memcheck, racecheck, synccheck of whole time steps on the cavity
# NOTE: compute-sanitizer is closed on this pool (profiles/r2_y_compute_sanitizer_closed_on_this_pool.txt holds its message); the script is kept for boxes where it runs.
# (marching kernels, TMA rings, cluster tail), the outflow channel (generic kernels + Krylov route) and the no-alloc entry point.
# usage: tools/sanitize.sh <tag>   ->  gpurun_out/<tag>_sanitizer_<tool>.log
tag=${1:-r2}
export NSGPU_GRAPH=0
cat > /tmp/san_prog.py <<'PY'
import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np
import navierstokessolver_b200 as ns
from navierstokessolver_b200 import capi
for n, nx in ((128, 128), (130, 96)):
    s = ns.GpuFluidSolver(ns.GridDesc.cavity(n, nx=nx), 0.25 / n, 1000.0)
    i = s.step(2); s.export_begin(1); s.step(1); s.export_wait()
    print("cavity", n, nx, "status", i.status, "it", i.it_u, i.it_v, i.it_phi)
    s.close()
g = ns.GridDesc(np.full(96, 1.0 / 48), np.full(48, 1.0 / 48), [-1, 0, 1, 0], [0, 1, 0, -1], [0, 2, 4, 2], [1.0, 0.0, 0.0, 0.0])
s = ns.GpuFluidSolver(g, 0.25 / 48, 100.0)
i = s.step(2)
print("channel status", i.status, "it", i.it_u, i.it_v, i.it_phi)
s.close()
print("noalloc", capi.minmax_noalloc(np.arange(1000.0), -np.arange(1000.0)))
PY
for tool in memcheck racecheck synccheck; do
  compute-sanitizer --tool $tool --print-limit 20 python /tmp/san_prog.py > gpurun_out/${tag}_sanitizer_${tool}.log 2>&1
  echo "== $tool: $(grep -c "=========.*\(Error\|error\|hazard\)" gpurun_out/${tag}_sanitizer_${tool}.log) report lines; $(grep "ERROR SUMMARY\|RACECHECK SUMMARY" gpurun_out/${tag}_sanitizer_${tool}.log | tail -1)"
  grep "cavity\|channel\|noalloc" gpurun_out/${tag}_sanitizer_${tool}.log | head -5
done
