"""Per-pass timing table on N ranks (NSGPU_TRACE=1, direct launches), launched by torch.distributed.run:
  python -m torch.distributed.run --nproc-per-node 2 tools/trace_step_mgpu.py [ny] [nx_per_rank] [steps]"""
import os, sys
os.environ["NSGPU_TRACE"] = "1"
os.environ["NSGPU_GRAPH"] = "0"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import navierstokessolver_b200 as ns
from navierstokessolver_b200 import capi
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
ny = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nxl = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
k = int(sys.argv[3]) if len(sys.argv) > 3 else 5
torch.cuda.set_device(local)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
box = [capi.nccl_unique_id() if rank == 0 else None]
dist.broadcast_object_list(box, src=0)
gnx = nxl * world
s = ns.GpuFluidSolver(ns.GridDesc.cavity(ny, nx=gnx, lx=gnx / float(ny)), 0.25 / ny, 1000.0, device=local, rank=rank, world=world, nccl_id=box[0])
try:   # (timing experiments with NSGPU_NOWAIT=1 produce wrong fields: a NaN status must not lose the trace table)
    s.step(3)
    info = s.step(k)
    if rank == 0:
        print("ranks", world, "steps", k, "ms/step", info.ms_total / k, "it", info.it_u, info.it_v, info.it_phi, "phases", s.phase_ms(), file=sys.stderr)
except ns.NsGpuError as e:
    print("rank", rank, "step failed:", e, file=sys.stderr)
s.close()
dist.barrier()
dist.destroy_process_group()
