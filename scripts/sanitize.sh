#!/bin/bash
# compute-sanitizer runs over a small, fast subset of the GPU tests (one process per tool), SURVEY.md section 5.2.
#   bash scripts/sanitize.sh [memcheck|racecheck|synccheck|initcheck ...]      (default: memcheck synccheck)
# Logs go to gpurun_out/sanitizer_<tool>.log; exit status is non-zero if any tool reports an error.
# The subset keeps every kernel family (tcgen05 conv / wgrad, elementwise, loss head, optimizer) at tiny shapes:
# the sanitizer slows kernels down by two to three orders of magnitude.
mkdir -p gpurun_out
tools="$@"
if [ -z "$tools" ]; then tools="memcheck synccheck"; fi
select="channel_sum or pack_input or adam or (wgrad_halo and cins0) or (conv3x3_halo and cins3) or dgrad_split_and_mask"
status=0
for tool in $tools; do
  log=gpurun_out/sanitizer_$tool.log
  echo "=== compute-sanitizer --tool $tool" | tee $log
  timeout ${SANITIZE_TIMEOUT:-600} compute-sanitizer --tool $tool --error-exitcode 86 --print-limit 20 \
      python -m pytest tests/test_gpu_elementwise.py tests/test_gpu_wgrad.py tests/test_gpu_conv.py -m gpu -q -x \
      -k "$select" --no-header -p no:cacheprovider >> $log 2>&1
  rc=$?
  tail -n 6 $log
  echo "=== $tool exit $rc" | tee -a $log
  if [ $rc -ne 0 ]; then status=1; fi
done
exit $status
