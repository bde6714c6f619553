#!/bin/bash
# usage: examples/FSDP2/run_example.sh [NPROC]      runs the FSDP2 variants and the FSDP2×TP example
set -euo pipefail
cd "$(dirname "$0")/../.."
source examples/set_env.sh
N=${1:-2}
TR="python -m torch.distributed.run --standalone --local-addr 127.0.0.1 --nproc-per-node $N"
$TR examples/FSDP2/fsdp2_main.py --steps 5
$TR examples/FSDP2/fsdp2_main.py --steps 5 --mixed-precision --explicit-prefetch --ac
$TR examples/FSDP2/fsdp2_main.py --steps 4 --ckpt-dir /tmp/fsdp2_ckpt && $TR examples/FSDP2/fsdp2_main.py --steps 6 --ckpt-dir /tmp/fsdp2_ckpt
if [ "$N" -ge 2 ] && [ $((N % 2)) -eq 0 ]; then $TR examples/FSDP2/fsdp2_tp_main.py --tp 2 --steps 5; fi
