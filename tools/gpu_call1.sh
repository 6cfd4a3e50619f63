#!/bin/bash
# call 1: parity tests with the marching sweep, then sweep / step timings for several chunk lengths
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/c1_pytest.log
tail -5 gpurun_out/c1_pytest.log
for ri in 0 16 32 64 128 512; do
  echo "== NSGPU_RI=$ri" ; NSGPU_RI=$ri timeout 300 python tools/quick_gpu.py 4096
done > gpurun_out/c1_quick.log 2>&1
echo "== NSGPU_MARCH=0" >> gpurun_out/c1_quick.log; NSGPU_MARCH=0 timeout 300 python tools/quick_gpu.py 4096 >> gpurun_out/c1_quick.log 2>&1
echo "== sizes" >> gpurun_out/c1_quick.log; timeout 600 python tools/quick_gpu.py 1024 2048 8192 16384 >> gpurun_out/c1_quick.log 2>&1
cat gpurun_out/c1_quick.log
