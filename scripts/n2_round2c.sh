#!/bin/bash
# 2-GPU check of the side-stream warm-up + capture: DP2 headline (ZeRO-1 path in the graph), EP2 MoE capture, TP2+SP capture and clean exit
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
P=29850
run() { name=$1; to=$2; shift 2; P=$((P+1)); timeout $to $TR --master-port $P bench.py --gpus 2 "$@" > gpurun_out/n2c_$name.log 2>&1; echo "$name rc=$?"; grep '^{' gpurun_out/n2c_$name.log | tail -1 | cut -c1-200; grep -o '"cuda_graph": [a-z]*' gpurun_out/n2c_$name.log | tail -1; }
run dp2_headline 200 --steps 20 --warmup 5
SCALETORCH_B200_GRAPH_MP=ep run moe_l8_ep2_graph 240 --steps 5 --warmup 3 --skip_e2e --model Qwen3-30B-A3B-L8 --ep 2 --micro_batch_size 1 --seq_len 4096
SCALETORCH_B200_GRAPH_MP=tp run q8b_tp2sp_graph 240 --steps 5 --warmup 3 --skip_e2e --model Qwen3-8B --tp 2 --sequence_parallel --micro_batch_size 1 --seq_len 4096
grep -h "Warning\|Error\|error" gpurun_out/n2c_*.log | cut -c1-250 | sort | uniq -c | sort -rn | head -8
