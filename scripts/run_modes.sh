#!/bin/bash
# Every single-process / pipeline mode BASELINE.json asks for, on the GPUs of one lease, with this build's kernels and the
# reference arm beside each row.  JSON lines go to gpurun_out/modes_<N>gpu.jsonl; stderr to gpurun_out/modes_<N>gpu.err.
#   bash scripts/run_modes.sh            # uses every visible GPU
#   MODES="dp mp" bash scripts/run_modes.sh
N=$(python -c "import torch; print(torch.cuda.device_count())")
OUT=gpurun_out/modes_${N}gpu.jsonl
ERR=gpurun_out/modes_${N}gpu.err
mkdir -p gpurun_out
: > $OUT; : > $ERR
MODES=${MODES:-"dp dpref mp mpref milesial"}
PORT=29800
run() { echo "### $*" >> $ERR; timeout ${TMO:-300} "$@" >> $OUT 2>> $ERR || echo "{\"failed\": \"$*\"}" >> $OUT; }
trun() { local n=$1; shift; PORT=$((PORT+1)); run python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT "$@"; }
has() { [[ " $MODES " == *" $1 "* ]]; }

if has dp; then
  for g in 2 4 8; do
    [ $g -le $N ] || continue
    CUDA_VISIBLE_DEVICES=$(seq -s, 0 $((g-1))) run python scripts/bench_modes.py --mode dp --batch-per-gpu 4          # weak: global 4*g
    CUDA_VISIBLE_DEVICES=$(seq -s, 0 $((g-1))) run python scripts/bench_modes.py --mode dp --batch 4                  # the reference's -b 4
    CUDA_VISIBLE_DEVICES=$(seq -s, 0 $((g-1))) run python scripts/bench_modes.py --mode dp --batch-per-gpu 4 --variant milesial
  done
fi
if has dpref; then
  for g in 2 4 8; do
    [ $g -le $N ] || continue
    CUDA_VISIBLE_DEVICES=$(seq -s, 0 $((g-1))) run python scripts/bench_modes.py --mode dp-ref --batch-per-gpu 4
  done
fi
if has mp; then
  trun 2 scripts/bench_modes.py --mode mp --micro 4 --batch 8 --schedule gpipe
  trun 2 scripts/bench_modes.py --mode mp --micro 4 --batch 8 --schedule 1f1b
  trun 2 scripts/bench_modes.py --mode mp --micro 8 --batch 8 --schedule 1f1b
  trun 2 scripts/bench_modes.py --mode mp --micro 2 --batch 4 --schedule gpipe
  trun 2 scripts/bench_modes.py --mode mp --micro 4 --batch 8 --schedule gpipe --reserve 0
  trun 2 scripts/bench_modes.py --mode mp --micro 4 --batch 8 --schedule gpipe --variant milesial
  if [ $N -ge 4 ]; then
    trun 4 scripts/bench_modes.py --mode mp --micro 8 --batch 8 --schedule gpipe
    trun 4 scripts/bench_modes.py --mode mp --micro 8 --batch 8 --schedule 1f1b
    trun 4 scripts/bench_modes.py --mode mp --micro 8 --batch 16 --schedule 1f1b
    trun 4 scripts/bench_modes.py --mode mp --micro 8 --batch 8 --schedule 1f1b --variant milesial
  fi
fi
if has mpsweep; then
  # SMs left to NCCL x NCCL channels x micro-batch size, 2 stages, 1F1B
  for rc in "8 4" "8 8" "16 8" "16 16" "24 12"; do
    set -- $rc
    trun 2 scripts/bench_modes.py --mode mp --micro 4 --batch 16 --schedule 1f1b --reserve $1 --nccl-channels $2
  done
  trun 2 scripts/bench_modes.py --mode mp --micro 8 --batch 32 --schedule 1f1b
  trun 2 scripts/bench_modes.py --mode mp --micro 4 --batch 16 --schedule gpipe
  trun 2 scripts/bench_modes.py --mode mp --micro 4 --batch 16 --schedule 1f1b --variant milesial
fi
if has mp4; then
  trun 4 scripts/bench_modes.py --mode mp --micro 8 --batch 8 --schedule 1f1b
  trun 4 scripts/bench_modes.py --mode mp --micro 8 --batch 16 --schedule 1f1b
  trun 4 scripts/bench_modes.py --mode mp --micro 8 --batch 32 --schedule 1f1b
  trun 4 scripts/bench_modes.py --mode mp --micro 8 --batch 32 --schedule gpipe
  trun 4 scripts/bench_modes.py --mode mp --micro 8 --batch 32 --schedule 1f1b --reserve 24 --nccl-channels 24
  trun 4 scripts/bench_modes.py --mode mp --micro 8 --batch 32 --schedule 1f1b --variant milesial
  trun 2 scripts/bench_modes.py --mode mp --micro 4 --batch 16 --schedule 1f1b --reserve 24 --nccl-channels 24
  trun 2 scripts/bench_modes.py --mode mp --micro 4 --batch 16 --schedule 1f1b --reserve 16 --nccl-channels 32
fi
if has dp4; then
  for g in 4; do
    CUDA_VISIBLE_DEVICES=$(seq -s, 0 $((g-1))) run python scripts/bench_modes.py --mode dp --batch-per-gpu 4
    CUDA_VISIBLE_DEVICES=$(seq -s, 0 $((g-1))) run python scripts/bench_modes.py --mode dp --batch 4
    CUDA_VISIBLE_DEVICES=$(seq -s, 0 $((g-1))) run python scripts/bench_modes.py --mode dp --batch-per-gpu 4 --variant milesial
    CUDA_VISIBLE_DEVICES=$(seq -s, 0 $((g-1))) run python scripts/bench_modes.py --mode dp-ref --batch-per-gpu 4
  done
  CUDA_VISIBLE_DEVICES=0,1 run python scripts/bench_modes.py --mode dp-ref --batch-per-gpu 4
fi
if has eight; then
  CUDA_VISIBLE_DEVICES=0,1,2,3,4,5,6,7 run python scripts/bench_modes.py --mode dp --batch-per-gpu 4
  CUDA_VISIBLE_DEVICES=0,1,2,3,4,5,6,7 run python scripts/bench_modes.py --mode dp --batch 4
  CUDA_VISIBLE_DEVICES=0,1,2,3,4,5,6,7 run python scripts/bench_modes.py --mode dp --batch-per-gpu 4 --variant milesial
  CUDA_VISIBLE_DEVICES=0,1,2,3,4,5,6,7 run python scripts/bench_modes.py --mode dp-ref --batch-per-gpu 4
  trun 8 bench.py --gpus 8 --model milesial --steps 10 --warmup 3
  trun 8 bench.py --gpus 8 --model milesial-bilinear --batch 16 --steps 10 --warmup 3
  trun 8 bench.py --gpus 8 --model milesial-bilinear --batch 16 --steps 10 --warmup 3 --torch-baseline
  trun 8 bench.py --gpus 8 --impl reference --amp bf16 --steps 10 --warmup 3
  trun 8 scripts/bench_collective.py --mb 31.04
  trun 8 scripts/bench_collective.py --mb 124.15
fi
if has mpref; then
  run python scripts/bench_modes.py --mode mp-ref --batch 4
  run python scripts/bench_modes.py --mode mp-ref --batch 8
fi
if has milesial; then
  # BASELINE.json config 5 and the north-star model under DDP, with the same module on torch bf16 autocast beside it
  for m in milesial milesial-bilinear; do
    trun $N bench.py --gpus $N --model $m --steps 10 --warmup 3
    trun $N bench.py --gpus $N --model $m --steps 10 --warmup 3 --torch-baseline
  done
  trun $N bench.py --gpus $N --model milesial-bilinear --batch 16 --steps 10 --warmup 3
  trun $N bench.py --gpus $N --model milesial-bilinear --batch 16 --steps 10 --warmup 3 --torch-baseline
fi
cat $OUT
