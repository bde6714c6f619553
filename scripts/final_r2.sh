#!/bin/bash
# Round-2 closing run on ONE GPU: regression, both bench arms, kernel tables, one ncu capture (after its plain run).
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/final_gpu_tests.log 2>&1
echo "pytest rc=$? $(tail -1 gpurun_out/final_gpu_tests.log)"
timeout 400 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/final_ref_n1.json 2> gpurun_out/final_ref_n1.err
echo "reference rc=$?"; cat gpurun_out/final_ref_n1.json | cut -c1-300
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/final_ours_n1.json 2> gpurun_out/final_ours_n1.err
echo "ours rc=$?"; cat gpurun_out/final_ours_n1.json | cut -c1-300
timeout 240 python tools/bench_kernels.py gemm > gpurun_out/gemm_v3_vs_cublas.jsonl 2> gpurun_out/bench_gemm.err; echo "gemm rc=$?"
timeout 200 python tools/bench_kernels.py attn > gpurun_out/attention_v5_vs_sdpa.jsonl 2> gpurun_out/bench_attn.err; echo "attn rc=$?"
timeout 200 python tools/bench_kernels.py mem > gpurun_out/membound_kernels_r2.jsonl 2> gpurun_out/bench_mem.err; echo "mem rc=$?"
timeout 200 python tools/debug_capture_moe_1gpu.py > gpurun_out/capture_moe_1gpu.log 2>&1; echo "capture probe rc=$?"; grep '^{' gpurun_out/capture_moe_1gpu.log | cut -c1-300
timeout 200 python tools/prof_kernels_once.py > gpurun_out/prof_once_plain.log 2>&1
rc=$?; echo "prof plain rc=$rc"; tail -3 gpurun_out/prof_once_plain.log
if [ $rc -eq 0 ]; then
  timeout 600 ncu --set full --clock-control none --import-source on \
    -k 'regex:^(gemm_|flash_attn|attn_bwd|grouped_|moe_router|rmsnorm|add_rmsnorm|swiglu|qknorm|ce_|adamw|sumsq)' -c 40 \
    -o gpurun_out/prof_kernels_r2 -f python tools/prof_kernels_once.py > gpurun_out/ncu_kernels_r2.log 2>&1
  echo "ncu rc=$?"; ls -la gpurun_out/prof_kernels_r2.ncu-rep
fi
