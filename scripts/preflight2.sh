#!/bin/bash
# 2-GPU pre-flight of the bench.py parallelism flags before the 8-GPU runs.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
run() { name=$1; shift; timeout 300 $TR --master-port $((29700 + RANDOM % 200)) bench.py --gpus 2 --steps 5 --warmup 3 --skip_e2e "$@" > gpurun_out/pre_$name.log 2>&1; echo "$name rc=$?"; tail -1 gpurun_out/pre_$name.log | cut -c1-400; }
run tp2sp --model Qwen3-8B --tp 2 --sequence_parallel --micro_batch_size 1 --seq_len 4096
run cp2 --model Qwen3-1.7B --cp 2 --micro_batch_size 1 --seq_len 8192
run pp2 --model Qwen3-1.7B --pp 2 --micro_batch_size 1 --seq_len 4096 --grad_accum 4
