#!/bin/bash
# 8-GPU record of the Qwen3-8B TP2-CP2-DP2 + SP target configuration with the step replayed from a CUDA graph
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29871 bench.py --gpus 8 \
  --steps 5 --warmup 3 --skip_e2e --model Qwen3-8B --tp 2 --cp 2 --sequence_parallel --micro_batch_size 1 --seq_len 4096 > gpurun_out/s8r2b_q8b_graph.log 2>&1
echo "rc=$?"; grep '^{' gpurun_out/s8r2b_q8b_graph.log | tail -1 | cut -c1-900; grep -h "Warning" gpurun_out/s8r2b_q8b_graph.log | cut -c1-250 | sort | uniq -c | head -5
