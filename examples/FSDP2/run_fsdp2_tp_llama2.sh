#!/bin/bash
# usage: examples/FSDP2/run_fsdp2_tp_llama2.sh [NPROC] [TP] [SIZE]
#   2-D parallel Llama-2: FSDP2 over the dp mesh dimension x tensor / sequence parallelism over tp (NPROC must be a multiple of TP).
set -euo pipefail
cd "$(dirname "$0")/../.."
source examples/set_env.sh
N=${1:-4}; TP=${2:-2}; SIZE=${3:-tiny}
if [ $((N % TP)) -ne 0 ]; then echo "NPROC ($N) must be a multiple of TP ($TP)"; exit 1; fi
python -m torch.distributed.run --standalone --local-addr 127.0.0.1 --nproc-per-node "$N" examples/FSDP2/fsdp2_tp_llama2_main.py \
    --tp "$TP" --size "$SIZE" --steps "${STEPS:-10}" --batch "${BATCH:-4}" --seq "${SEQ:-128}"
