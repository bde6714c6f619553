#!/bin/bash
# 2-GPU MoE check: Qwen3-30B-A3B shapes (8 layers), EP2: fused peer-memory path vs c10d path vs reference arm.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
A="--gpus 2 --steps 5 --warmup 3 --model Qwen3-30B-A3B-L8 --ep 2 --micro_batch_size 1 --seq_len 4096 --skip_e2e"
timeout 300 $TR --master-port 29611 bench.py $A > gpurun_out/moe_ep2_fused.log 2>&1; echo "fused rc=$?"; tail -1 gpurun_out/moe_ep2_fused.log
SCALETORCH_B200_FUSED_COMM=0 timeout 300 $TR --master-port 29612 bench.py $A > gpurun_out/moe_ep2_nccl.log 2>&1; echo "nccl rc=$?"; tail -1 gpurun_out/moe_ep2_nccl.log
timeout 400 $TR --master-port 29613 bench.py $A --impl reference > gpurun_out/moe_ep2_ref.log 2>&1; echo "ref rc=$?"; tail -3 gpurun_out/moe_ep2_ref.log
