#!/bin/bash
# Run every GPU test file in its own process (a CUDA fault in one must not hide the others); logs -> gpurun_out/
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
for f in ${@:-tests/test_gpu_*.py}; do
  n=$(basename $f .py)
  timeout 600 python -m pytest $f -q -m gpu -x --timeout 300 > gpurun_out/$n.log 2>&1
  echo "== $n rc=$? $(tail -1 gpurun_out/$n.log)"
done
