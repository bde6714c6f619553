#!/bin/bash
# 2-GPU model-level records of round 2 (MoE EP2, CP2 long context, TP2+SP fused vs c10d); JSON lines in gpurun_out/n2_*.log
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
P=29700
run() { name=$1; to=$2; shift 2; P=$((P+1)); timeout $to $TR --master-port $P bench.py --gpus 2 "$@" > gpurun_out/n2_$name.log 2>&1; echo "$name rc=$?"; tail -1 gpurun_out/n2_$name.log | cut -c1-420; }
run moe_l8_ep2 240 --steps 5 --warmup 3 --skip_e2e --model Qwen3-30B-A3B-L8 --ep 2 --micro_batch_size 1 --seq_len 4096
run q17b_cp2 240 --steps 5 --warmup 3 --skip_e2e --model Qwen3-1.7B --cp 2 --micro_batch_size 1 --seq_len 8192
run q8b_tp2sp_fused 300 --steps 5 --warmup 3 --skip_e2e --model Qwen3-8B --tp 2 --sequence_parallel --micro_batch_size 1 --seq_len 4096
SCALETORCH_B200_FUSED_COMM=0 run q8b_tp2sp_c10d 300 --steps 5 --warmup 3 --skip_e2e --model Qwen3-8B --tp 2 --sequence_parallel --micro_batch_size 1 --seq_len 4096
