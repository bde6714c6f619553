#!/bin/bash
# DP gradient all-reduce sweep at N GPUs (default 2): CTA budget x transport (NVLS multimem vs two-shot peer loads vs NCCL).
N=${1:-2}
mkdir -p gpurun_out; rm -f gpurun_out/dp_sweep.log
run() { echo "== $1" >> gpurun_out/dp_sweep.log; env $1 SCALETORCH_B200_VERBOSE=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $N --steps 10 --warmup 4 --skip_e2e 2>&1 | grep -E "^\{|GradWindow|Error" | sed -E 's/.*"ms_per_step": ([0-9.]+).*/ms_per_step \1/' >> gpurun_out/dp_sweep.log; }
run "SCALETORCH_B200_DP_SM_BUDGET=8" 29711
run "SCALETORCH_B200_DP_SM_BUDGET=16" 29712
run "SCALETORCH_B200_DP_SM_BUDGET=32" 29713
run "SCALETORCH_B200_DP_SM_BUDGET=64" 29714
run "SCALETORCH_B200_DP_SM_BUDGET=148" 29715
run "SCALETORCH_B200_DP_SM_BUDGET=148 SCALETORCH_B200_NO_MULTIMEM=1" 29716
run "SCALETORCH_B200_FUSED_COMM=0" 29717
cat gpurun_out/dp_sweep.log
