#!/bin/bash
# 8-GPU records of round 2 for the two target configurations of BASELINE.json (ours arm; JSON lines in gpurun_out/s8r2_*.log)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
P=29800
run() { name=$1; to=$2; shift 2; P=$((P+1)); timeout $to $TR --master-port $P bench.py --gpus 8 "$@" > gpurun_out/s8r2_$name.log 2>&1; echo "$name rc=$?"; tail -1 gpurun_out/s8r2_$name.log | cut -c1-900; }
run q8b_tp2cp2dp2_s4096 200 --steps 5 --warmup 3 --skip_e2e --model Qwen3-8B --tp 2 --cp 2 --sequence_parallel --micro_batch_size 1 --seq_len 4096
SCALETORCH_B200_EP_CAPACITY=2.0 run moe30b_ep8_s4096 280 --steps 5 --warmup 3 --skip_e2e --model Qwen3-30B-A3B --ep 8 --micro_batch_size 1 --seq_len 4096
