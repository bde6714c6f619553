#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
A="--gpus 2 --steps 5 --warmup 3 --skip_e2e --model Qwen3-1.7B --cp 2 --micro_batch_size 1 --seq_len 8192"
timeout 500 python -m pytest tests/test_multigpu_model_gpu.py -x -q -m gpu -k "cp" --timeout 200 --timeout-method=thread -p no:cacheprovider 2>&1 | tail -5
timeout 300 $TR --master-port 29801 bench.py $A > gpurun_out/cp2_zz.log 2>&1; echo "zigzag rc=$?"; tail -1 gpurun_out/cp2_zz.log | cut -c1-300
timeout 300 $TR --master-port 29802 bench.py $A --cp_zigzag 0 > gpurun_out/cp2_contig.log 2>&1; echo "contiguous rc=$?"; tail -1 gpurun_out/cp2_contig.log | cut -c1-300
