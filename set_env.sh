#!/bin/bash
# Environment of a B200 training node (source set_env.sh).  The reference's root set_env.sh loads the Ascend CANN
# toolkit; the counterpart here is CUDA 12.8+ on PATH plus the few NCCL / PyTorch switches the framework assumes.
export CUDA_HOME=${CUDA_HOME:-/usr/local/cuda}
export PATH="$CUDA_HOME/bin:$PATH"
export LD_LIBRARY_PATH="$CUDA_HOME/lib64:${LD_LIBRARY_PATH:-}"
export MASTER_ADDR=${MASTER_ADDR:-127.0.0.1}          # the container hostname may not resolve
export MASTER_PORT=${MASTER_PORT:-29500}
export OMP_NUM_THREADS=${OMP_NUM_THREADS:-8}
export CUDA_DEVICE_MAX_CONNECTIONS=${CUDA_DEVICE_MAX_CONNECTIONS:-8}   # several hardware queues: the comm stream must not queue behind compute
export NCCL_NVLS_ENABLE=${NCCL_NVLS_ENABLE:-1}         # in-switch reductions for the c10d fallback / baseline paths
export TORCH_NCCL_AVOID_RECORD_STREAMS=1
export TORCH_NCCL_ASYNC_ERROR_HANDLING=1
export PYTHONPATH="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd):${PYTHONPATH:-}"
# build the native library once per checkout (nvcc, sm_100a; ~25 s):  python -c "import __graft_entry__ as g; g.build()"
