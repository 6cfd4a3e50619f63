"""phi cycles per step on long cavities (one GPU): python tools/aspect_cycles.py  -> JSON lines"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import navierstokessolver_b200 as ns
for ny, nx in ((512, 512), (512, 1024), (512, 4096), (1024, 8192), (256, 4096)):
    grid = ns.GridDesc.cavity(ny, nx=nx, lx=nx / float(ny))
    s = ns.GpuFluidSolver(grid, 0.25 / ny, 1000.0)
    its = []
    for k in range(4):
        info = s.step(1)
        its.append([info.it_u, info.it_v, info.it_phi])
    print(json.dumps({"ny": ny, "nx": nx, "levels": s.num_levels, "iterations": its, "status": info.status}))
    s.close()
