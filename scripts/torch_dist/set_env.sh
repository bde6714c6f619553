#!/bin/bash
# Environment defaults for one 8×B200 NVSwitch node (reference: scripts/torch_dist/set_env.sh sets HCCL/Ascend vars).
# Source it:  source scripts/torch_dist/set_env.sh
export MASTER_ADDR=${MASTER_ADDR:-127.0.0.1}
export MASTER_PORT=${MASTER_PORT:-29500}
export OMP_NUM_THREADS=${OMP_NUM_THREADS:-8}
export TOKENIZERS_PARALLELISM=false
export CUDA_DEVICE_MAX_CONNECTIONS=${CUDA_DEVICE_MAX_CONNECTIONS:-32}   # compute + comm + copy streams need their own HW queues
export NCCL_NVLS_ENABLE=${NCCL_NVLS_ENABLE:-1}                          # in-switch reductions when the fabric exposes multicast
export NCCL_DEBUG=${NCCL_DEBUG:-WARN}
export TORCH_NCCL_ASYNC_ERROR_HANDLING=1
export TORCH_NCCL_AVOID_RECORD_STREAMS=1
export PYTORCH_CUDA_ALLOC_CONF=${PYTORCH_CUDA_ALLOC_CONF:-expandable_segments:True}
# framework switches (scaletorch_b200/env.py)
export SCALETORCH_B200_FUSED_COMM=${SCALETORCH_B200_FUSED_COMM:-1}      # peer-memory fused collectives (0 = c10d/NCCL path)
export SCALETORCH_B200_SYMM_WINDOW_MB=${SCALETORCH_B200_SYMM_WINDOW_MB:-1024}
export SCALETORCH_B200_EP_CAPACITY=${SCALETORCH_B200_EP_CAPACITY:-2.0}
