mkdir -p gpurun_out; rm -f gpurun_out/dp_sweep.log
run() { echo "== $1" >> gpurun_out/dp_sweep.log; env $1 SCALETORCH_B200_VERBOSE=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $2 bench.py --gpus 2 --steps 6 --warmup 3 --skip_e2e 2>&1 | grep -E "^\{|GradWindow|Error" | sed -E 's/.*"ms_per_step": ([0-9.]+).*/ms_per_step \1/' >> gpurun_out/dp_sweep.log; }
run "SCALETORCH_B200_DP_SM_BUDGET=16" 29711
run "SCALETORCH_B200_DP_SM_BUDGET=64" 29712
run "SCALETORCH_B200_DP_SM_BUDGET=148" 29713
run "SCALETORCH_B200_DP_SM_BUDGET=32 SCALETORCH_B200_NO_MULTIMEM=1" 29714
run "SCALETORCH_B200_DP_SM_BUDGET=32 SCALETORCH_B200_DP_BUCKET=67108864" 29715
run "SCALETORCH_B200_FUSED_COMM=0 SCALETORCH_B200_DP_BUCKET=67108864" 29716
cat gpurun_out/dp_sweep.log
