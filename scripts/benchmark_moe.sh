#!/bin/bash
# Qwen3-30B-A3B expert-parallel layouts on 8 GPUs (reference: scripts/benchmark_moe.sh exercises EP2-TP4 and EP4-TP2;
# EP8 hangs there).  SCALETORCH_B200_FUSED_COMM=0 selects the c10d all-to-all path for an A/B comparison.
set -uo pipefail
cd "$(dirname "$0")/.."
source scripts/torch_dist/set_env.sh
mkdir -p results
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
i=0
for LAYOUT in "--ep 8" "--ep 4 --tp 2" "--ep 2 --tp 4" "--ep 2 --tp 4 --sequence_parallel"; do
  for SEQ in 2048 4096; do
    i=$((i + 1))
    $TR --master-port $((29500 + i)) bench.py --gpus 8 --steps "${STEPS:-5}" --warmup 3 --skip_e2e --model Qwen3-30B-A3B $LAYOUT \
        --micro_batch_size 1 --seq_len $SEQ 2> "results/moe_${i}.err" | tail -1 | tee -a results/benchmark_moe.jsonl
  done
done
