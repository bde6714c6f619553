#!/bin/bash
# Runs every DeviceMesh / DTensor demo on N processes (default 4; CPU/gloo when no GPU is visible).
set -euo pipefail
cd "$(dirname "$0")/../.."
source examples/set_env.sh
N=${1:-4}
for f in device_mesh_api dtensor_demo tensor_parallel_demo sequence_parallel_demo fsdp_tp_demo fsdp_dp_demo manual_process_group; do
  echo "=== $f ==="
  python -m torch.distributed.run --standalone --local-addr 127.0.0.1 --nproc-per-node "$N" examples/device_mesh/$f.py
done
