#!/bin/bash
# device-side multi-rank solves: parity vs single GPU (graph and direct), 2-rank bench, single-GPU regression
mkdir -p gpurun_out
run2() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
for c in "256 128 5" "200 36 3" "130 64 3" "1024 512 3"; do
  run2 29511 tools/mgpu_check.py $c 2>&1 | tail -2
done
echo "--- direct launches"
NSGPU_GRAPH=0 run2 29513 tools/mgpu_check.py 256 128 5 2>&1 | tail -2
echo "--- bench 2 ranks"
run2 29512 bench.py --gpus 2 --steps 10 --warmup 3 2>gpurun_out/c31_b2.err | tail -1 > gpurun_out/c31_bench2.json; head -c 1800 gpurun_out/c31_bench2.json; tail -5 gpurun_out/c31_b2.err
echo "--- single"
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 10 --warmup 3 2>/dev/null | head -c 900
