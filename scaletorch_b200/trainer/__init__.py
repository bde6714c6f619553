from .config import (CheckpointArguments, DataArguments, LoggingArguments, LrSchedulerArguments, ModelArguments,
                     OptimizerArguments, ParallelArguments, ScaleTorchArguments, TrainingArguments, validate_config)
from .evaluation import PeriodicEvaluator, evaluate_loss, reduce_eval_stats
from .graphed import GraphedTrainStep
from .dist_setup import (cleanup_distributed, cleanup_distributed_training, get_dtype,
                         initialize_distributed_training)
from .lr_scheduler import create_lr_scheduler, register_scheduler
from .metrics import AsyncScalarReader, log_training_metrics
from .model_builder import create_model, create_optimizer, get_tensor_shapes
from .train_step import clip_gradients, train_step, training_step

__all__ = [
    "AsyncScalarReader", "CheckpointArguments", "GraphedTrainStep", "PeriodicEvaluator", "evaluate_loss", "reduce_eval_stats", "DataArguments", "LoggingArguments", "LrSchedulerArguments", "ModelArguments",
    "OptimizerArguments", "ParallelArguments", "ScaleTorchArguments", "TrainingArguments", "cleanup_distributed", "cleanup_distributed_training",
    "clip_gradients", "create_lr_scheduler", "create_model", "create_optimizer", "get_dtype", "get_tensor_shapes",
    "initialize_distributed_training", "log_training_metrics", "register_scheduler", "train_step", "training_step", "validate_config",
]
