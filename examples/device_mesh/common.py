"""Process-group bring-up shared by the DeviceMesh / DTensor demos (NCCL on GPUs, gloo on CPU)."""
import os

import torch
import torch.distributed as dist


def setup():
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29512")
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    if torch.cuda.is_available():
        dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
        torch.cuda.set_device(dev)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    else:
        dev = torch.device("cpu")
        dist.init_process_group("gloo", rank=rank, world_size=world)
    return rank, world, dev


def ordered_print(msg: str) -> None:
    for r in range(dist.get_world_size()):
        if r == dist.get_rank():
            print(f"[rank {r}] {msg}", flush=True)
        dist.barrier()


def finish() -> None:
    dist.barrier()
    dist.destroy_process_group()
