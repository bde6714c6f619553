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

# Llama-2 variants (reference: run_wikitext2.sh, run_fsdp2_tp_llama2.sh)
#   torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 examples/FSDP2/fsdp2_llama2_main.py --size tiny --steps 20 --ckpt-dir /tmp/llama2_ckpt --save-every 10
#   torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 4 examples/FSDP2/fsdp2_tp_llama2_main.py --tp 2 --size tiny --steps 10
#   python examples/FSDP2/test_smoke_mp.py
