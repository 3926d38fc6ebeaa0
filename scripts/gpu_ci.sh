#!/bin/bash
# Runs every GPU test file in its own process (a trapped kernel poisons the CUDA context of its process only),
# each under a timeout, logs into gpurun_out/.  Usage on the GPU box:  bash scripts/gpu_ci.sh [files...]
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu_info.txt 2>&1
files="$@"
if [ -z "$files" ]; then files=$(ls tests/test_gpu_*.py); fi
status=0
for f in $files; do
  name=$(basename $f .py)
  echo "=== $f" | tee -a gpurun_out/ci_summary.txt
  timeout 600 python -m pytest $f -m gpu -q --no-header -p no:cacheprovider > gpurun_out/$name.log 2>&1
  rc=$?
  tail -n 25 gpurun_out/$name.log | tee -a gpurun_out/ci_summary.txt
  echo "=== $f exit $rc" | tee -a gpurun_out/ci_summary.txt
  if [ $rc -ne 0 ]; then status=1; fi
done
exit $status
