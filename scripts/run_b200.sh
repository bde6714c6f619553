#!/bin/bash
# Preset launcher (reference: scripts/run_npu.sh:40-56 -- modes max_mfu / max_speed / balanced / min_mem map to
# (micro batch, sequence, GA, GC) presets, data-parallel over all local GPUs).
#   usage: scripts/run_b200.sh [mode] [model] [extra tools/train.py flags...]
# Measured on one B200 (Qwen3-0.6B, bf16, profiles/bench_r1_*.json): BS2 S2048 -> 127.8K tok/s, 32.06 ms/step.
set -euo pipefail
cd "$(dirname "$0")/.."
MODE=${1:-balanced}; MODEL=${2:-Qwen3-0.6B}; shift $(( $# > 2 ? 2 : $# ))
source scripts/torch_dist/set_env.sh
case "$MODE" in
  max_mfu)   MBS=1; SEQ=16384; GA=1; GC=false ;;   # long sequences: attention FLOPs dominate, best tensor-core utilisation
  max_speed) MBS=8; SEQ=2048;  GA=1; GC=false ;;   # most tokens/s
  balanced)  MBS=2; SEQ=4096;  GA=2; GC=false ;;
  min_mem)   MBS=1; SEQ=2048;  GA=4; GC=true  ;;
  *) echo "mode must be one of max_mfu|max_speed|balanced|min_mem"; exit 2 ;;
esac
NGPU=${NGPU:-$(nvidia-smi -L | wc -l)}
python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NGPU" --master-addr 127.0.0.1 --master-port "${MASTER_PORT:-29500}" \
  tools/train.py --model_name_or_path "$MODEL" --synthetic_data true --data_parallel_size "$NGPU" \
  --micro_batch_size $MBS --sequence_length $SEQ --gradient_accumulation_steps $GA --gradient_checkpointing $GC \
  --total_train_steps "${STEPS:-50}" "$@"
