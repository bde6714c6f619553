#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
SCALETORCH_B200_VERBOSE=1 timeout 200 python tools/debug_capture_moe_1gpu.py step > gpurun_out/capture_moe_verbose.log 2>&1; echo "verbose rc=$?"
grep "flat_params\|^{" gpurun_out/capture_moe_verbose.log | cut -c1-250
SCALETORCH_B200_GRAPH_SIDE_STREAM=1 timeout 200 python tools/debug_capture_moe_1gpu.py step > gpurun_out/capture_moe_side.log 2>&1; echo "side rc=$?"
grep "^{\|Warning" gpurun_out/capture_moe_side.log | cut -c1-250
SCALETORCH_B200_GRAPH_SIDE_STREAM=1 timeout 300 python -m pytest tests/test_graph_gpu.py -x -q -p no:cacheprovider 2>&1 | tail -2
