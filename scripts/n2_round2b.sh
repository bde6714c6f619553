#!/bin/bash
# 2-GPU follow-up: TP2+SP with the measured fused-op policy, CUDA-graph capture attempts for TP and EP (bounded by timeout)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
P=29750
run() { name=$1; to=$2; shift 2; P=$((P+1)); timeout $to $TR --master-port $P bench.py --gpus 2 "$@" > gpurun_out/n2_$name.log 2>&1; echo "$name rc=$?"; tail -1 gpurun_out/n2_$name.log | cut -c1-330; }
run q8b_tp2sp_auto 300 --steps 5 --warmup 3 --skip_e2e --model Qwen3-8B --tp 2 --sequence_parallel --micro_batch_size 1 --seq_len 4096
SCALETORCH_B200_GRAPH_MP=tp run q8b_tp2sp_graph 240 --steps 5 --warmup 3 --skip_e2e --model Qwen3-8B --tp 2 --sequence_parallel --micro_batch_size 1 --seq_len 4096
SCALETORCH_B200_GRAPH_MP=ep run moe_l8_ep2_graph 240 --steps 5 --warmup 3 --skip_e2e --model Qwen3-30B-A3B-L8 --ep 2 --micro_batch_size 1 --seq_len 4096
grep -h "cuda_graph\|Warning\|warn" gpurun_out/n2_q8b_tp2sp_graph.log gpurun_out/n2_moe_l8_ep2_graph.log | cut -c1-300 | tail -6
