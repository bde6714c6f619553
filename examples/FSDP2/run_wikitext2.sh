#!/bin/bash
# usage: examples/FSDP2/run_wikitext2.sh [NPROC] [SIZE] [DATA_KIND] [PATH]
#   FSDP2 training of the Llama-2-shaped model on a token stream, with periodic DCP checkpoints and resume.
#   DATA_KIND: synthetic (default, no files needed) | tokens (flat .npy / .bin of token ids) | text (plain text file, byte-level ids)
#   e.g.  examples/FSDP2/run_wikitext2.sh 8 7b tokens /data/wikitext2_train.npy
set -euo pipefail
cd "$(dirname "$0")/../.."
source examples/set_env.sh
N=${1:-2}; SIZE=${2:-tiny}; KIND=${3:-synthetic}; DATA=${4:-}
CKPT=${CKPT_DIR:-/tmp/fsdp2_llama2_ckpt}
ARGS="--size $SIZE --steps ${STEPS:-40} --batch ${BATCH:-4} --seq ${SEQ:-128} --warmup 5 --data $KIND --ckpt-dir $CKPT --save-every ${SAVE_EVERY:-20}"
[ -n "$DATA" ] && ARGS="$ARGS --path $DATA"
[ "${BF16:-0}" = "1" ] && ARGS="$ARGS --bf16"
python -m torch.distributed.run --standalone --local-addr 127.0.0.1 --nproc-per-node "$N" examples/FSDP2/fsdp2_llama2_main.py $ARGS
echo "re-run the same command to resume from $CKPT/latest"
