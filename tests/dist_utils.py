"""Spawn helpers for the multi-process CPU (gloo) tests."""
import os
import socket
import traceback

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, fn, args):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fn(rank, world, *args)
    except Exception:
        traceback.print_exc()
        raise
    finally:
        try:
            from scaletorch_b200.parallel import process_group as pg

            pg.process_group_manager._instance = None
        except Exception:
            pass
        dist.destroy_process_group()


_START_METHOD = os.environ.get("STB_TEST_START_METHOD", "forkserver")
if _START_METHOD == "forkserver":
    # Children are forked from a clean, single-threaded server process that has already imported the heavy modules, so
    # a rank starts in milliseconds instead of re-importing torch + transformers (several seconds per rank, ~250 ranks
    # per suite run).  The server is started lazily by the first test and lives as long as the pytest process.
    import multiprocessing as _pymp

    try:
        _pymp.set_forkserver_preload(["torch", "torch.distributed", "numpy", "transformers", "scaletorch_b200",
                                      "scaletorch_b200.trainer", "scaletorch_b200.models"])
    except Exception:  # pragma: no cover - platform without forkserver
        _START_METHOD = "spawn"


def run_distributed(fn, world_size: int, *args) -> None:
    """Runs ``fn(rank, world_size, *args)`` on ``world_size`` CPU processes with a gloo process group."""
    global _START_METHOD
    try:
        mp.start_processes(_worker, args=(world_size, _free_port(), fn, args), nprocs=world_size, join=True,
                           start_method=_START_METHOD)
    except (OSError, ValueError, ImportError):
        if _START_METHOD == "spawn":
            raise
        _START_METHOD = "spawn"               # the fork server could not be started here: plain spawn from now on
        mp.start_processes(_worker, args=(world_size, _free_port(), fn, args), nprocs=world_size, join=True,
                           start_method="spawn")


class TinyCfg:
    """Attribute bag in the shape of a HF config."""

    def __init__(self, **kw):
        base = dict(vocab_size=256, hidden_size=64, num_attention_heads=4, num_key_value_heads=2, head_dim=16,
                    max_position_embeddings=128, num_hidden_layers=2, intermediate_size=128, rms_norm_eps=1e-6,
                    rope_theta=10000.0, tie_word_embeddings=False, model_type="llama", num_experts=4,
                    num_experts_per_tok=2, moe_intermediate_size=32, norm_topk_prob=True, router_aux_loss_coef=0.01)
        base.update(kw)
        self.__dict__.update(base)


def rel_err(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()
