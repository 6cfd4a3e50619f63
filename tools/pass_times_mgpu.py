"""Fine-level pass times on N ranks (nsgpu_time_passes: CUDA events around back-to-back launches), launched by torch.distributed.run:
  python -m torch.distributed.run --nproc-per-node 2 tools/pass_times_mgpu.py [ny] [nx_per_rank] [n]
kinds: 0 plain sweep, 1 first pass of a phi cycle, 2 last pass (prolongation + |r|^2 summed over the ranks), 3 SOR double sweep"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import navierstokessolver_b200 as ns
from navierstokessolver_b200 import capi
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
ny = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nxl = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
n = int(sys.argv[3]) if len(sys.argv) > 3 else 50
torch.cuda.set_device(local)
box = [None]
if world > 1:
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    box = [capi.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
gnx = nxl * world
s = ns.GpuFluidSolver(ns.GridDesc.cavity(ny, nx=gnx, lx=gnx / float(ny)), 0.25 / ny, 1000.0, device=local, rank=rank, world=world, nccl_id=box[0])
out = {}
try:
    s.step(3)
    info = s.step(5)
    out["ms/step"] = round(info.ms_total / 5, 3)
    out["it"] = (info.it_u, info.it_v, info.it_phi)
    out["phases"] = [round(x, 3) for x in s.phase_ms()]
except ns.NsGpuError as e:
    out["step error"] = str(e)[:80]
try:
    last = int(os.environ.get("PASS_LAST", "0"))   # the kind timed last: its CTA stamps are the ones NSGPU_XRED_DBG=2 prints
    for kind in [k for k in (1, 2, 3, 0) if k != last] + [last]:
        s.time_passes(kind, 5)
        if os.environ.get("PASS_PARITY"):   # odd counts swap the roles of the ping-pong arrays: does the time depend on which array is read?
            out[f"kind{kind}_us"] = [round(s.time_passes(kind, m) * 1000.0 / m, 1) for m in (n, n + 1, n, n + 1, n + 1, n)]
        else:
            out[f"kind{kind}_us"] = round(min(s.time_passes(kind, n) for _ in range(3)) * 1000.0 / n, 1)
except ns.NsGpuError as e:
    out["error"] = str(e)
if rank == 0:
    print("ranks", world, {k: v for k, v in os.environ.items() if k.startswith("NSGPU_")}, out, flush=True)
s.close()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
