#!/bin/bash
# One call on a 4-GPU box: DP / MP rows of BASELINE.md for both arms, plus the 4-GPU DDP bench.  Everything under
# `timeout`; results in gpurun_out/modes4.jsonl.
out=gpurun_out/modes4.jsonl; : > $out
run() { timeout 300 "$@" >> $out 2>> gpurun_out/modes4.err || echo "{\"failed\": \"$*\"}" >> $out; }
run python scripts/bench_modes.py --mode dp
run python scripts/bench_modes.py --mode dp --batch 4
run python scripts/bench_modes.py --mode dp-ref
run python scripts/bench_modes.py --mode dp-ref --batch 4
CUDA_VISIBLE_DEVICES=0,1 run python scripts/bench_modes.py --mode mp-ref
run python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29801 scripts/bench_modes.py --mode mp --micro 2 --batch 4
run python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29802 scripts/bench_modes.py --mode mp --micro 4 --batch 8
run python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29803 scripts/bench_modes.py --mode mp --micro 8 --batch 8
run python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29804 scripts/bench_modes.py --mode mp --micro 8 --batch 8 --schedule 1f1b
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29805 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/bench_native_4.json 2> gpurun_out/bench_native_4.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29806 bench.py --impl reference --gpus 4 --steps 10 --warmup 3 > gpurun_out/bench_ref_4.json 2> gpurun_out/bench_ref_4.err
cat $out
