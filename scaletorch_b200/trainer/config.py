"""Training configuration dataclasses, parsed with ``transformers.HfArgumentParser``.

Same eight argument groups, field names and defaults as the reference
(``scaletorch/trainer/config.py:31-475``) so existing launch scripts keep working; every field is a
``--flag``.  Additions (all optional): ``--sequence_parallel``, ``--ep_replicate_data``,
``--fused_comm``, ``--grad_wire_dtype``, ``--synthetic_data``, ``--model_config_path``,
``--aux_loss``.  Validation fixes: ``expert_parallel_size`` is checked, and ``nccl``/``gloo``
are the only back-ends (Ascend ``hccl`` is dropped).
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field

_VALID_LR_SCHEDULERS = {"linear", "cosine", "polynomial", "step", "onecycle", "constant"}
_VALID_OPTIMIZERS = {"adamw", "adam", "sgd", "lamb"}


@dataclass
class DataArguments:
    """Dataset and loader flags: dataset name or path, tokenizer, subset, split, worker counts, sample cap, pinned memory, synthetic data."""
    data_path: str = field(default="./data", metadata={"help": "Dataset download directory"})
    dataset_name: str = field(default="wikitext2", metadata={"help": "Dataset name or local directory"})
    tokenizer_name_or_path: str = field(default="facebook/opt-125m", metadata={"help": "Tokenizer name or path"})
    subset_name: str | None = field(default=None, metadata={"help": "Subset name"})
    split: str = field(default="train", metadata={"help": "Split name"})
    num_proc: int = field(default=1, metadata={"help": "Number of processes for data processing"})
    num_workers: int = field(default=2, metadata={"help": "Number of DataLoader workers"})
    num_samples: int | None = field(default=None, metadata={"help": "Number of samples to use (None = all)"})
    pin_memory: bool = field(default=True, metadata={"help": "Whether to pin host memory"})
    synthetic_data: bool = field(default=False, metadata={"help": "Random tokens of the configured shape (benchmarks; no dataset / tokenizer needed)"})


@dataclass
class ModelArguments:
    """Which model to build (name or path of an HF config / checkpoint), optional overrides of layer / head counts, attention switch, dtype."""
    model_name_or_path: str = field(default="facebook/opt-125m", metadata={"help": "Model name, local HF directory or a name under configs/"})
    num_hidden_layers: int | None = field(default=None, metadata={"help": "Override the number of hidden layers"})
    num_attention_heads: int | None = field(default=None, metadata={"help": "Override the number of attention heads"})
    num_key_value_heads: int | None = field(default=None, metadata={"help": "Override the number of key-value heads"})
    use_flash_attention: bool = field(default=True, metadata={"help": "Use the flash attention kernel (False = eager oracle)"})
    dtype: str = field(default="bfloat16", metadata={"help": "Training dtype: bfloat16 | float32"})

    def __post_init__(self) -> None:
        cfg = resolve_model_config(self.model_name_or_path, quiet=True)
        if cfg is not None:
            for name in ("num_hidden_layers", "num_attention_heads", "num_key_value_heads"):
                if getattr(self, name) is None:
                    setattr(self, name, getattr(cfg, name, None))


@dataclass
class ParallelArguments:
    """Sizes of the five parallel axes, the pipeline schedule, the backend, and the switches of the parallel implementations (sequence parallel, zig-zag CP, EP capacity and data replication, fused collectives, gradient wire dtype)."""
    tensor_parallel_size: int = field(default=1, metadata={"help": "Size of tensor parallelism"})
    pipeline_parallel_size: int = field(default=1, metadata={"help": "Size of pipeline parallelism"})
    data_parallel_size: int = field(default=1, metadata={"help": "Size of data parallelism"})
    context_parallel_size: int = field(default=1, metadata={"help": "Size of context parallelism"})
    expert_parallel_size: int = field(default=1, metadata={"help": "Size of expert parallelism"})
    pipeline_parallel_engine: str = field(default="1f1b", metadata={"help": "Pipeline schedule: 1f1b | afab"})
    backend: str = field(default="nccl", metadata={"help": "Distributed backend: nccl | gloo"})
    sequence_parallel: bool = field(default=False, metadata={"help": "Megatron sequence parallelism over the TP group (also enabled by env SEQUENCE_PARALLEL=1)"})
    ep_capacity_factor: float | None = field(default=None, metadata={"help": "Receive capacity of the peer-memory EP path as a multiple of tokens*top_k. Default (None, and env SCALETORCH_B200_EP_CAPACITY unset) is dropless (= ep size); a smaller value bounds memory and DROPS the pairs that do not fit (counted in ep_dropped_pairs)"})
    cp_zigzag: bool = field(default=True, metadata={"help": "Load-balanced (zig-zag) causal context parallelism: rank r owns chunks r and 2cp-1-r (needs S divisible by 2cp; falls back to the contiguous split otherwise)"})
    ep_replicate_data: bool = field(default=False, metadata={"help": "Reference semantics: every EP rank of a dp_rank reads the same samples"})
    fused_comm: bool = field(default=True, metadata={"help": "Use the fused compute+collective kernels over NVLink peer memory when available"})
    grad_wire_dtype: str = field(default="float32", metadata={"help": "dtype of DP gradient buckets on the wire: float32 | bfloat16"})

    def __post_init__(self) -> None:
        for name in ("data_parallel_size", "tensor_parallel_size", "pipeline_parallel_size", "context_parallel_size",
                     "expert_parallel_size"):
            if getattr(self, name) < 1:
                raise ValueError(f"{name} must be >= 1, got {getattr(self, name)}")
        if self.pipeline_parallel_engine not in {"1f1b", "afab"}:
            raise ValueError(f'pipeline_parallel_engine must be "1f1b" or "afab", got {self.pipeline_parallel_engine}')
        if self.backend not in {"nccl", "gloo"}:
            raise ValueError(f"backend must be one of {{nccl, gloo}}, got {self.backend}")
        if self.grad_wire_dtype not in {"float32", "bfloat16"}:
            raise ValueError(f"grad_wire_dtype must be float32 or bfloat16, got {self.grad_wire_dtype}")
        if os.environ.get("SEQUENCE_PARALLEL", "0") == "1":
            self.sequence_parallel = True


@dataclass
class LrSchedulerArguments:
    """Learning-rate schedule: type, warm-up steps and the parameters of each decay family; validated in ``__post_init__``."""
    lr_scheduler_type: str = field(default="linear", metadata={"help": "Type of learning rate scheduler"})
    warmup_steps: int = field(default=0, metadata={"help": "Number of warmup steps"})
    T_max: int | None = field(default=None, metadata={"help": "Maximum number of iterations for cosine annealing"})
    eta_min: float = field(default=0.0, metadata={"help": "Minimum learning rate for cosine annealing"})
    power: float = field(default=1.0, metadata={"help": "Power factor for polynomial decay"})
    step_size: int = field(default=1, metadata={"help": "Step size for step decay"})
    gamma: float = field(default=0.1, metadata={"help": "Multiplicative factor for step decay"})
    max_lr: float | None = field(default=None, metadata={"help": "Upper learning rate boundary for OneCycleLR"})
    pct_start: float = field(default=0.3, metadata={"help": "Fraction of the cycle spent increasing LR (OneCycleLR)"})

    def __post_init__(self) -> None:
        if self.lr_scheduler_type not in _VALID_LR_SCHEDULERS:
            from .lr_scheduler import registered_schedulers

            if self.lr_scheduler_type not in registered_schedulers():
                raise ValueError(f"lr_scheduler_type must be one of {sorted(_VALID_LR_SCHEDULERS)}, got {self.lr_scheduler_type}")
        if self.warmup_steps < 0:
            raise ValueError(f"warmup_steps must be >= 0, got {self.warmup_steps}")
        t = self.lr_scheduler_type
        if t == "cosine" and self.eta_min < 0:
            raise ValueError(f"eta_min must be >= 0, got {self.eta_min}")
        if t == "polynomial" and self.power <= 0:
            raise ValueError(f"power must be > 0, got {self.power}")
        if t == "step":
            if self.step_size <= 0:
                raise ValueError(f"step_size must be > 0, got {self.step_size}")
            if self.gamma <= 0 or self.gamma > 1:
                raise ValueError(f"gamma must be in (0, 1], got {self.gamma}")
        if t == "onecycle" and not 0 < self.pct_start < 1:
            raise ValueError(f"pct_start must be in (0, 1), got {self.pct_start}")


@dataclass
class OptimizerArguments:
    """Optimiser type, learning rate, betas, weight decay, fused-kernel switch."""
    optimizer_type: str = field(default="adamw", metadata={"help": "Type of optimizer"})
    weight_decay: float = field(default=0.0, metadata={"help": "Weight decay"})
    use_fused_adam: bool = field(default=True, metadata={"help": "Use the fused AdamW kernel over the flat parameter store"})
    betas: list[float] = field(default_factory=lambda: [0.9, 0.999], metadata={"help": "Betas for Adam (two floats: --betas 0.9 0.95)"})
    learning_rate: float = field(default=1e-3, metadata={"help": "Learning rate"})

    def __post_init__(self) -> None:
        if self.optimizer_type not in _VALID_OPTIMIZERS:
            raise ValueError(f"optimizer_type must be one of {sorted(_VALID_OPTIMIZERS)}, got {self.optimizer_type}")
        self.betas = tuple(float(b) for b in self.betas)   # parsed as a list (nargs="+"), used as a 2-tuple
        if len(self.betas) != 2:
            raise ValueError(f"betas must be two floats, got {self.betas}")


@dataclass
class TrainingArguments:
    """Batch geometry (micro-batch, sequence length, accumulation), stop criteria, seed, clipping, gradient checkpointing, ZeRO stage, native loader, CUDA-graph capture, lagged logging, periodic evaluation, CPU run, MoE aux loss."""
    batch_size: int = field(default=64, metadata={"help": "Training batch size"})
    test_batch_size: int = field(default=1000, metadata={"help": "Evaluation batch size"})
    micro_batch_size: int | None = field(default=None, metadata={"help": "Micro batch size"})
    gradient_accumulation_steps: int = field(default=1, metadata={"help": "Number of gradient accumulation steps"})
    gradient_checkpointing: bool = field(default=False, metadata={"help": "Recompute decoder layers in backward"})
    max_grad_norm: float | None = field(default=1.0, metadata={"help": "Max gradient norm (None disables clipping)"})
    epochs: int = field(default=5, metadata={"help": "Number of training epochs"})
    seed: int = field(default=1, metadata={"help": "Random seed"})
    sequence_length: int | None = field(default=1024, metadata={"help": "Sequence length"})
    log_interval: int = field(default=100, metadata={"help": "Logging frequency"})
    max_tokens: int | None = field(default=None, metadata={"help": "Stop after this many trained tokens"})
    native_loader: bool = field(default=False, metadata={"help": "Produce micro-batches with the C++ worker threads (data/native_loader.py); synthetic or flat-token data"})
    token_file: str | None = field(default=None, metadata={"help": "Flat int32 token file (.npy / raw .bin) for --native_loader"})
    zero_stage: int = field(default=0, metadata={"help": "1 = shard AdamW state and update over the gradient-sync group (trainer/zero.py)"})
    use_cuda_graph: bool = field(default=False, metadata={"help": "Replay the whole step from a CUDA graph (trainer/graphed.py)"})
    log_lag: int = field(default=1, metadata={"help": "Log step i after enqueueing step i+lag (0 = synchronous logging)"})
    nonfinite_loss: str = field(default="abort", metadata={"help": "What to do when the logged loss is NaN / Inf: 'abort' (exit non-zero so that a supervisor resumes from the last checkpoint) or 'warn' (keep going, as the reference does)"})
    eval_interval: int = field(default=0, metadata={"help": "Evaluate the held-out loss every N optimiser steps (0 = never; with pipeline stages the pass runs forward-only through the pipeline)"})
    eval_batches: int = field(default=8, metadata={"help": "Micro-batches per data rank in one evaluation (the same batches every time)"})
    eval_token_file: str | None = field(default=None, metadata={"help": "Flat int32 token file (.npy / raw .bin) with the held-out tokens; default: --eval_split of the dataset, or synthetic data with another seed"})
    eval_split: str = field(default="validation", metadata={"help": "Dataset split used for evaluation when no --eval_token_file is given"})
    total_train_steps: int | None = field(default=None, metadata={"help": "Stop after this many steps"})
    use_cpu: bool = field(default=False, metadata={"help": "Run on CPU with the gloo backend"})
    aux_loss: bool = field(default=True, metadata={"help": "Add the MoE load-balancing loss to the training loss"})

    def __post_init__(self) -> None:
        if self.gradient_accumulation_steps < 1:
            raise ValueError(f"gradient_accumulation_steps must be >= 1, got {self.gradient_accumulation_steps}")
        if self.micro_batch_size is not None and self.micro_batch_size < 1:
            raise ValueError(f"micro_batch_size must be >= 1, got {self.micro_batch_size}")
        if self.sequence_length is not None and self.sequence_length < 1:
            raise ValueError(f"sequence_length must be >= 1, got {self.sequence_length}")
        if self.nonfinite_loss not in ("abort", "warn"):
            raise ValueError(f"nonfinite_loss must be 'abort' or 'warn', got {self.nonfinite_loss!r}")
        if self.eval_interval < 0:
            raise ValueError(f"eval_interval must be >= 0, got {self.eval_interval}")
        if self.eval_batches < 1:
            raise ValueError(f"eval_batches must be >= 1, got {self.eval_batches}")


@dataclass
class CheckpointArguments:
    """Where and how often to save, what to resume from, whether to load pretrained weights, background writing and save-on-signal."""
    work_dir: str = field(default="./work_dir", metadata={"help": "Directory to save checkpoints"})
    save_model_checkpoint: bool = field(default=True, metadata={"help": "Whether to save model checkpoints"})
    save_frequency: int = field(default=300, metadata={"help": "Checkpoint frequency in steps (0 = never)"})
    resume_path: str = field(default="", metadata={"help": "Checkpoint directory to resume from; 'latest' = the newest complete checkpoint under --work_dir (none: start from scratch)"})
    resume_reshard: bool = field(default=False, metadata={"help": "Allow --resume_path to point at a checkpoint written under another TP / PP / EP layout: the weights are merged and re-cut for this layout, step / token counters and the data position carry over, the optimiser moments restart from zero"})
    keep_last_checkpoints: int = field(default=0, metadata={"help": "Keep only the N newest complete checkpoints under --work_dir, delete older ones after each save (0 = keep all)"})
    save_on_signal: bool = field(default=False, metadata={"help": "On SIGTERM / SIGUSR1 (preemption) finish the current step, let all ranks agree, write a checkpoint and exit 0"})
    signal_check_interval: int = field(default=10, metadata={"help": "With --save_on_signal: ranks compare their stop flags every this many steps (one 8-byte gloo all-reduce)"})
    async_checkpoint: bool = field(default=False, metadata={"help": "Write checkpoints on a background thread: the step only pays for the copy of the state to host memory"})
    load_pretrained_weights: bool = field(default=False, metadata={"help": "Load and KEEP the HF safetensors weights found in model_name_or_path"})

    def __post_init__(self) -> None:
        if self.save_frequency < 0:
            raise ValueError(f"save_frequency must be >= 0, got {self.save_frequency}")


@dataclass
class LoggingArguments:
    """wandb switches and names, the profiler window and trace directory, NVTX ranges, the step watchdog timeout."""
    use_wandb: bool = field(default=False, metadata={"help": "Log to Weights & Biases"})
    project_name: str = field(default="scaletorch", metadata={"help": "Project name for logging"})
    experiment_name: str | None = field(default=None, metadata={"help": "Experiment name for logging"})
    profile_start_step: int | None = field(default=None, metadata={"help": "First step of the torch.profiler window"})
    profile_end_step: int | None = field(default=None, metadata={"help": "Step at which the profiler window closes"})
    trace_dir: str = field(default="traces", metadata={"help": "Where Chrome traces / kernel tables are written"})
    nvtx: bool = field(default=False, metadata={"help": "Emit NVTX ranges (step / phase) for Nsight Systems"})
    step_timeout_s: float = field(default=0.0, metadata={"help": "Watchdog: abort if a step takes longer (0 = off)"})


@dataclass
class ScaleTorchArguments(DataArguments, ModelArguments, ParallelArguments, LrSchedulerArguments, OptimizerArguments,
                          TrainingArguments, CheckpointArguments, LoggingArguments):
    """All training arguments (multiple inheritance of the eight groups, like the reference)."""

    global_batch_size: int = field(init=False, default=0)
    global_batch_size_token: int | None = field(init=False, default=None)

    def __post_init__(self) -> None:
        for cls in (DataArguments, ModelArguments, ParallelArguments, LrSchedulerArguments, OptimizerArguments,
                    TrainingArguments, CheckpointArguments, LoggingArguments):
            post = cls.__dict__.get("__post_init__")
            if post is not None:
                post(self)
        if self.micro_batch_size is None:
            self.micro_batch_size = self.batch_size
        if self.sequence_length is not None and self.sequence_length % self.context_parallel_size != 0:
            raise ValueError(f"sequence_length ({self.sequence_length}) must be divisible by context_parallel_size "
                             f"({self.context_parallel_size})")
        data_replicas = self.data_parallel_size * (1 if self.ep_replicate_data else self.expert_parallel_size)
        self.global_batch_size = data_replicas * self.micro_batch_size * self.gradient_accumulation_steps
        if self.sequence_length is not None:
            self.global_batch_size_token = self.global_batch_size * self.sequence_length

    def validate_world_size(self, world_size: int) -> None:
        expected = (self.tensor_parallel_size * self.pipeline_parallel_size * self.data_parallel_size *
                    self.context_parallel_size * self.expert_parallel_size)
        if world_size != expected:
            raise ValueError(
                f"world_size ({world_size}) != TP ({self.tensor_parallel_size}) * PP ({self.pipeline_parallel_size}) * "
                f"DP ({self.data_parallel_size}) * CP ({self.context_parallel_size}) * EP ({self.expert_parallel_size}) "
                f"= {expected}")


def validate_config(config: ScaleTorchArguments, world_size: int) -> None:
    """Check that the parallel sizes multiply to ``world_size`` (raises ``ValueError`` otherwise)."""
    config.validate_world_size(world_size)


# --------------------------------------------------------------------------- model config resolution
_CONFIG_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "configs")
_PACKAGED_CONFIG_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "configs")   # pip-installed copy


# keys a config.json must spell out for the fast path below; anything less goes through ``AutoConfig`` so that the
# family's class defaults fill the gaps exactly as HuggingFace would
_DENSE_KEYS = ("hidden_size", "intermediate_size", "num_hidden_layers", "num_attention_heads", "num_key_value_heads",
               "vocab_size", "max_position_embeddings", "rms_norm_eps", "tie_word_embeddings", "rope_theta")
_COMPLETE_CONFIG_KEYS = {
    "llama": _DENSE_KEYS,
    "qwen3": _DENSE_KEYS + ("head_dim",),
    "qwen3_moe": _DENSE_KEYS + ("head_dim", "num_experts", "num_experts_per_tok", "moe_intermediate_size", "norm_topk_prob",
                                "decoder_sparse_step", "mlp_only_layers", "router_aux_loss_coef"),
}


class LocalModelConfig:
    """The fields of a ``config.json`` as attributes (plus ``to_dict`` / ``get``), for model families this package builds.

    ``AutoConfig.from_pretrained`` imports the family's modelling package and with it torchvision, scikit-learn and
    ``torch._dynamo`` -- about ten seconds per process, on every rank of every launch.  The models here only read plain
    fields, so a config file that spells out every field the family needs is used as it is."""

    def __init__(self, values: dict, source: str):
        self.__dict__.update(values)
        self._name_or_path = source

    def to_dict(self) -> dict:
        return {k: v for k, v in self.__dict__.items() if not k.startswith("_")}

    def get(self, key: str, default=None):
        return self.__dict__.get(key, default)

    def __repr__(self) -> str:
        return f"LocalModelConfig({self._name_or_path}: {self.to_dict()})"


def _complete_local_config(directory: str):
    import json

    with open(os.path.join(directory, "config.json")) as f:
        values = json.load(f)
    need = _COMPLETE_CONFIG_KEYS.get(values.get("model_type"))
    if need is None or any(k not in values for k in need):
        return None
    return LocalModelConfig(values, directory)


def resolve_model_config(name_or_path: str, quiet: bool = False):
    """Returns the model configuration for a local HF directory, a vendored config under ``configs/<name>/config.json`` (the
    GPU box has no network) or, last, the HF hub name: a :class:`LocalModelConfig` when the file spells out every field of
    its family (all vendored configs do), otherwise the ``transformers.PretrainedConfig`` that ``AutoConfig`` builds
    (``SCALETORCH_B200_HF_AUTOCONFIG=1`` forces that path)."""
    base = os.path.basename(name_or_path.rstrip("/"))
    candidates = [name_or_path] + [os.path.join(d, n) for d in (_CONFIG_DIR, _PACKAGED_CONFIG_DIR) for n in (name_or_path, base)]
    for c in candidates:
        if os.path.isfile(os.path.join(c, "config.json")):
            if os.environ.get("SCALETORCH_B200_HF_AUTOCONFIG", "0") != "1":
                cfg = _complete_local_config(c)
                if cfg is not None:
                    return cfg
            from transformers import AutoConfig

            return AutoConfig.from_pretrained(c)
    try:
        from transformers import AutoConfig

        return AutoConfig.from_pretrained(name_or_path, local_files_only=True)
    except Exception as e:
        if not quiet:
            raise RuntimeError(f"cannot resolve a model config for '{name_or_path}': {e}") from e
        return None


def main() -> None:  # pragma: no cover - CLI helper
    """Parse the command line into ``ScaleTorchArguments`` and print the result (``python -m scaletorch_b200.trainer.config --help`` lists every flag)."""
    from transformers import HfArgumentParser

    (cfg,) = HfArgumentParser(ScaleTorchArguments).parse_args_into_dataclasses()
    print(cfg)


if __name__ == "__main__":  # pragma: no cover
    main()
