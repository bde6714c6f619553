#!/bin/bash
# Single-GPU MFU sweep over (micro batch, sequence length, gradient checkpointing) (reference: scripts/sweep_mfu.sh).
set -uo pipefail
cd "$(dirname "$0")/.."
MODEL=${1:-Qwen3-0.6B}
mkdir -p results
for CFG in "2 2048" "4 2048" "8 2048" "2 4096" "1 8192" "1 16384" "4 4096 gc" "2 8192 gc" "1 32768 gc"; do
  set -- $CFG
  GC=""; [ "${3:-}" = "gc" ] && GC="--gradient_checkpointing"
  python bench.py --model "$MODEL" --micro_batch_size "$1" --seq_len "$2" $GC --steps 10 --warmup 3 --skip_e2e 2>/dev/null | tail -1 \
    | python -c "import sys, json; d = json.loads(sys.stdin.read()); print(f\"BS$1 S$2 ${3:-}: {d['value']:,.0f} tok/s  {d['ms_per_step']:.1f} ms  MFU {d['mfu_pct_of_nominal_2.25PF']:.1f} % nominal / {d.get('mfu_pct_of_measured_sustained', 0):.1f} % of measured\")" \
    | tee -a "results/sweep_mfu_${MODEL}.txt"
done
