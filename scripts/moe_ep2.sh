#!/bin/bash
# 2-GPU MoE check: Qwen3-30B-A3B shapes (8 layers), EP2 with whole-step CUDA-graph replay
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
A="--gpus 2 --steps 5 --warmup 3 --model Qwen3-30B-A3B-L8 --ep 2 --micro_batch_size 1 --seq_len 4096 --skip_e2e"
timeout 300 $TR --master-port 29611 bench.py $A > gpurun_out/moe_ep2_graph.log 2>&1; echo "graph rc=$?"; tail -1 gpurun_out/moe_ep2_graph.log | cut -c1-250; grep -i "warn.*capture\|Error" gpurun_out/moe_ep2_graph.log | head -5
