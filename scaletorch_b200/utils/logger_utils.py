"""Rank-aware colourised logging (reference: ``scaletorch/utils/logger_utils.py:18-149``).

Rank 0 logs at the requested level to stdout; every other rank only surfaces errors.
"""
from __future__ import annotations

import logging
import os
import sys

_COLORS = {"DEBUG": "\033[36m", "INFO": "\033[32m", "WARNING": "\033[33m", "ERROR": "\033[31m", "CRITICAL": "\033[35m"}
_RESET = "\033[0m"
_loggers: dict[str, logging.Logger] = {}


def _get_distributed_rank() -> int:
    try:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            return dist.get_rank()
    except Exception:
        pass
    return int(os.environ.get("RANK", "0"))


class ColorfulFormatter(logging.Formatter):
    """Formatter that stamps each record with the distributed rank and colours the line by level when writing to a terminal."""
    def __init__(self, *args, use_color: bool = True, **kwargs):
        super().__init__(*args, **kwargs)
        self.use_color = use_color

    def format(self, record: logging.LogRecord) -> str:
        record.rank = _get_distributed_rank()
        msg = super().format(record)
        if self.use_color and record.levelname in _COLORS:
            return f"{_COLORS[record.levelname]}{msg}{_RESET}"
        return msg


def get_logger(name: str = "scaletorch_b200", log_file: str | None = None, log_level: int = logging.INFO,
               file_mode: str = "a") -> logging.Logger:
    """Rank-aware logger (cached by name): rank 0 logs to stdout (and ``log_file``) at ``log_level``, other ranks only errors to stderr."""
    if name in _loggers:
        return _loggers[name]
    logger = logging.getLogger(name)
    logger.propagate = False
    rank = _get_distributed_rank()
    fmt = "[%(asctime)s] [rank %(rank)s] [%(levelname)s] %(name)s: %(message)s"
    if rank == 0:
        h = logging.StreamHandler(sys.stdout)
        h.setFormatter(ColorfulFormatter(fmt, use_color=sys.stdout.isatty()))
        logger.addHandler(h)
        if log_file:
            os.makedirs(os.path.dirname(os.path.abspath(log_file)), exist_ok=True)
            fh = logging.FileHandler(log_file, file_mode)
            fh.setFormatter(ColorfulFormatter(fmt, use_color=False))
            logger.addHandler(fh)
        logger.setLevel(log_level)
    else:
        h = logging.StreamHandler(sys.stderr)
        h.setFormatter(ColorfulFormatter(fmt, use_color=False))
        logger.addHandler(h)
        logger.setLevel(logging.ERROR)
    _loggers[name] = logger
    return logger


def get_outdir(path: str, *paths: str, inc: bool = False) -> str:
    """Creates (and optionally auto-increments) an output directory."""
    outdir = os.path.join(path, *paths)
    if not os.path.exists(outdir):
        os.makedirs(outdir)
    elif inc:
        count = 1
        while os.path.exists(f"{outdir}-{count}"):
            count += 1
            if count >= 100:
                raise RuntimeError("too many output directories")
        outdir = f"{outdir}-{count}"
        os.makedirs(outdir)
    return outdir
