#!/bin/bash
# Environment for the examples (single node).  source examples/set_env.sh
export MASTER_ADDR=${MASTER_ADDR:-127.0.0.1}
export MASTER_PORT=${MASTER_PORT:-29511}
export OMP_NUM_THREADS=${OMP_NUM_THREADS:-4}
export TORCH_NCCL_AVOID_RECORD_STREAMS=1
export PYTHONPATH="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd):${PYTHONPATH:-}"
