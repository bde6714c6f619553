#!/bin/bash
# usage: examples/torch_examples/imagenet/run.sh [NPROC] [IMAGENET_DIR]
#   DDP ResNet-style training loop; without IMAGENET_DIR it trains on synthetic 3x64x64 images (a smoke / throughput check).
set -euo pipefail
cd "$(dirname "$0")/../../.."
source examples/set_env.sh
N=${1:-2}; DATA=${2:-}
ARGS="--iters ${ITERS:-20} --batch-size ${BATCH:-32}"
[ -n "$DATA" ] && ARGS="$ARGS --data $DATA"
python -m torch.distributed.run --standalone --local-addr 127.0.0.1 --nproc-per-node "$N" examples/torch_examples/imagenet/ddp_imagenet.py $ARGS
