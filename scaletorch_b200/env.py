"""Environment-variable names and defaults (reference: ``scaletorch/env.py:8-29``).

The variables are accepted as *inputs* for CLI parity with the reference; they are resolved once
into an explicit :class:`scaletorch_b200.parallel.context.ParallelContext` instead of being read
deep inside model constructors (SURVEY §7.4 D1 -- fixes the ordering bug of reference
``trainer/model_builder.py:67-88``).
"""
import os

ENV_FLASH_ATTENTION = "FLASH_ATTEN"  # legacy spelling kept
ENV_CONTEXT_PARALLEL = "CONTEXT_PARALLEL"
ENV_SEQUENCE_PARALLEL = "SEQUENCE_PARALLEL"
ENV_VERBOSE = "VERBOSE"
ENV_DTYPE = "DTYPE"
ENV_RANK = "RANK"
ENV_LOCAL_RANK = "LOCAL_RANK"
ENV_WORLD_SIZE = "WORLD_SIZE"
ENV_MASTER_ADDR = "MASTER_ADDR"
ENV_MASTER_PORT = "MASTER_PORT"
ENV_PRINT_LOCK = "SCALETORCH_PRINT_LOCK"
ENV_DEVICE_FLOPS = "SCALETORCH_DEVICE_FLOPS"
ENV_FORCE_TORCH = "SCALETORCH_B200_FORCE_TORCH"  # run the PyTorch oracle ops even on CUDA (debug)
ENV_FUSED_COMM = "SCALETORCH_B200_FUSED_COMM"    # 0 disables the symmetric-memory fused collectives

DEFAULT_DTYPE = "bfloat16"
DEFAULT_FLASH_ATTENTION_ENABLED = "1"
DEFAULT_PARALLELISM_DISABLED = "0"


def env_flag(name: str, default: str = "0") -> bool:
    """True when environment variable ``name`` (default ``default``) is ``1``."""
    return os.environ.get(name, default) == "1"


def env_int(name: str, default: int) -> int:
    """Integer value of environment variable ``name``, ``default`` when unset."""
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default
