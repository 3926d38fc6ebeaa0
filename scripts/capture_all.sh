#!/bin/bash
# ncu --set full of EVERY kernel launch of one training step (reference UNet and the BatchNorm UNet, batch 4, 640x960) on
# ONE GPU, each only after the same command has exited 0 without ncu.  Reports land in gpurun_out/; read them here with
#   python scripts/ncu_table.py gpurun_out/ncu_raw_ref.csv gpurun_out/labels_ref.json > profiles/ncu_r2_step_ref.md
# (the fused all-reduce kernel needs several ranks and is therefore not profiled with ncu: CUDA-event timings in
# profiles/collective_*.jsonl).
mkdir -p gpurun_out
for v in ${VARIANTS:-ref milesial}; do
  CMD="python scripts/one_step.py --variant $v --labels gpurun_out/labels_$v.json"
  $CMD > gpurun_out/plain_$v.log 2>&1 &&
  timeout ${NCU_TMO:-900} ncu --set full --clock-control none --import-source on --nvtx --nvtx-include "capture/" \
      -f -o gpurun_out/ncu_step_$v $CMD > gpurun_out/ncu_$v.log 2>&1
  echo "$v: rc=$? $(tail -1 gpurun_out/plain_$v.log)"
  # the report of a whole step is ~180 MB: export the raw page + the source pages that are read later, drop the report
  python scripts/ncu_export.py gpurun_out/ncu_step_$v.ncu-rep gpurun_out/labels_$v.json $v
done
du -sh gpurun_out
