#!/bin/bash
# Runs the DDP / FSDP examples on N processes (default 2; CPU/gloo when no GPU is visible).
set -euo pipefail
cd "$(dirname "$0")/../.."
source examples/set_env.sh
N=${1:-2}
TR="python -m torch.distributed.run --standalone --local-addr 127.0.0.1 --nproc-per-node $N"
$TR examples/torch_examples/mnist/ddp_mnist.py --epochs 2
$TR examples/torch_examples/mnist/fsdp_mnist.py --epochs 1
rm -f /tmp/minigpt_snapshot.pt
$TR examples/torch_examples/minigpt/ddp_minigpt.py --steps 20 && $TR examples/torch_examples/minigpt/ddp_minigpt.py --steps 40   # second run resumes at 20
$TR examples/torch_examples/imagenet/ddp_imagenet.py --iters 10
