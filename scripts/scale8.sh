#!/bin/bash
# 8-GPU record runs (one box): headline DP8 bench both arms, then the 3-D / EP / PP configs of BASELINE.md.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
P=29900
run() { name=$1; to=$2; shift 2; P=$((P+1)); timeout $to $TR --master-port $P bench.py --gpus 8 "$@" > gpurun_out/s8_$name.log 2>&1; echo "$name rc=$?"; tail -1 gpurun_out/s8_$name.log | cut -c1-330; }
run dp8_ours 150 --steps 20 --warmup 5
run dp8_ref 200 --steps 20 --warmup 5 --impl reference
run moe_ep8 240 --steps 5 --warmup 3 --skip_e2e --model Qwen3-30B-A3B --ep 8 --micro_batch_size 1 --seq_len 4096
run q8b_tp2cp2dp2 200 --steps 5 --warmup 3 --skip_e2e --model Qwen3-8B --tp 2 --cp 2 --sequence_parallel --micro_batch_size 1 --seq_len 8192
run q14b_tp4pp2 200 --steps 5 --warmup 3 --skip_e2e --model Qwen3-14B --tp 4 --pp 2 --sequence_parallel --micro_batch_size 1 --seq_len 4096 --grad_accum 8
run q8b_tp2cp2dp2_ref 240 --steps 5 --warmup 3 --skip_e2e --model Qwen3-8B --tp 2 --cp 2 --sequence_parallel --micro_batch_size 1 --seq_len 8192 --impl reference
nvidia-smi topo -m > gpurun_out/s8_topo.txt 2>&1
