"""Launched by torch.distributed.run with N ranks: N x-slabs of one cavity against the single-GPU run of the same
problem (rank 0 runs it afterwards on its own GPU).  Prints one JSON line; exit code 1 on mismatch.
  python -m torch.distributed.run --nproc-per-node 2 tools/mgpu_check.py [ny] [nx_per_rank] [steps] [cavity|channel]
"channel": inlet (u = 1) on the left, walls, outflow (bcType 4) on the right -- the generic kernels, NCCL halos and the augmented
phi solve of outflow domains on several ranks."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist
import navierstokessolver_b200 as ns
from navierstokessolver_b200 import capi

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
ny = int(sys.argv[1]) if len(sys.argv) > 1 else 256
nxl = int(sys.argv[2]) if len(sys.argv) > 2 else 128
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
kind = sys.argv[4] if len(sys.argv) > 4 else "cavity"
torch.cuda.set_device(local)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
box = [capi.nccl_unique_id() if rank == 0 else None]
dist.broadcast_object_list(box, src=0)
gnx = nxl * world
if kind == "channel":
    grid = ns.GridDesc(np.full(gnx, 1.0 / ny), np.full(ny, 1.0 / ny), [-1, 0, 1, 0], [0, 1, 0, -1], [0, 2, 4, 2], [1.0, 0.0, 0.0, 0.0])
    dt, re, rtol = 0.25 / ny, 100.0, 1e-10
else:
    grid = ns.GridDesc.cavity(ny, nx=gnx, lx=gnx / float(ny))
    dt, re, rtol = 0.25 / ny, 1000.0, 1e-10
s = ns.GpuFluidSolver(grid, dt, re, device=local, rank=rank, world=world, nccl_id=box[0], rtol_u=rtol, rtol_phi=rtol)
info = s.step(steps)
mine = {k: torch.from_numpy(s.get(k)).cuda() for k in (ns.U, ns.V, ns.PHI)}
out = {}
for k, t in mine.items():
    parts = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
    dist.gather(t, parts, dst=0)
    if rank == 0:
        out[k] = torch.cat(parts).cpu().numpy()
s.close()
ok = True
if rank == 0:
    one = ns.GpuFluidSolver(grid, dt, re, device=local, rtol_u=rtol, rtol_phi=rtol)
    i1 = one.step(steps)
    gaps = {}
    for k, name in ((ns.U, "u"), (ns.V, "v"), (ns.PHI, "phi")):
        ref = one.get(k)
        gaps[name] = float(np.abs(out[k] - ref).max() / np.abs(ref).max())
    # cavity: multigrid on both sides, 5e-8.  channel: Krylov iterations, whose paths separate with the summation order, and a v that
    # is small next to u -- the tolerance of the outflow whole-step tests (1e-6) applies
    ok = all(g < (1e-6 if kind == "channel" else 5e-8) for g in gaps.values()) and info.status == 0
    print(json.dumps({"world": world, "kind": kind, "grid": [gnx, ny], "steps": steps, "gaps_vs_single_gpu": gaps, "status": info.status,
                      "it": [info.it_u, info.it_v, info.it_phi], "it_single": [i1.it_u, i1.it_v, i1.it_phi],
                      "ms_per_step": info.ms_total / steps, "ms_per_step_single": i1.ms_total / steps, "ok": ok}), flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
