#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu_model_gpu.py -x -q -m gpu -k "dp" --timeout 200 --timeout-method=thread -p no:cacheprovider 2>&1 | tail -6
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29901 bench.py --gpus 2 --steps 20 --warmup 5 2>gpurun_out/n2_graph.err | tail -1 > gpurun_out/n2_graph.json; python -c "
import json; d=json.load(open('gpurun_out/n2_graph.json')); print('rowx', d['ms_per_step'], d['value'], d.get('cuda_graph'), d['e2e'])"
tail -3 gpurun_out/n2_graph.err
SCALETORCH_B200_EMB_ROW_EXCHANGE=0 timeout 200 $TR --master-port 29902 bench.py --gpus 2 --steps 20 --warmup 5 --skip_e2e 2>/dev/null | tail -1 | cut -c1-200
timeout 100 python bench.py --steps 20 --warmup 5 2>/dev/null | tail -1 > gpurun_out/bench_n1_graph.json; cut -c1-200 gpurun_out/bench_n1_graph.json
