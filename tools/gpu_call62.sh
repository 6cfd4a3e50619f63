#!/bin/bash
# 8 ranks, final: parity, weak scaling (bench default), strong scaling 16384^2
mkdir -p gpurun_out
run8() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
run8 29511 tools/mgpu_check.py 512 128 3 2>&1 | tail -1 | cut -c1-400
b() { run8 29512 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline "${@:2}" 2>gpurun_out/c62_$1.err | tail -1 > gpurun_out/c62_$1.json; python - $1 <<'PY'
import json,sys; d=json.load(open("gpurun_out/c62_%s.json"%sys.argv[1])); print(sys.argv[1], d["ms_per_step"], d["value"], d["phase_ms_last_step"], d["solver_iterations_last_step"], d["roofline"]["achieved"], d["config"]["levels"])
PY
}
b weak
b strong16k --ny 16384 --nx-per-gpu 2048
