#!/bin/bash
mkdir -p gpurun_out
R="timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512"
for a in "256 128 5" "1024 512 3" "200 36 3" "130 64 3"; do
  $R tools/mgpu_check.py $a > gpurun_out/c19.log 2>&1; grep '"world"' gpurun_out/c19.log || tail -8 gpurun_out/c19.log
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/c19_bench2.json 2> gpurun_out/c19.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/c19_bench2.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','solver_iterations_last_step','fine_poisson_sweeps_per_step','poisson_gcell_updates_per_s','host_wall_ms_per_step','gpu_launches')}, d['roofline']['frac'], d['e2e']['ms_per_step'])
PY
