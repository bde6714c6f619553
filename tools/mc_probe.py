import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as sm
rank=int(os.environ["RANK"]); world=int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda",rank))
t=sm.empty(1<<22, dtype=torch.uint8, device=torch.device("cuda",rank))
h=sm.rendezvous(t, group=dist.group.WORLD)
if rank==0:
    print("attrs", [a for a in dir(h) if not a.startswith("__")])
    try: print("multicast_ptr", h.multicast_ptr)
    except Exception as e: print("multicast_ptr err", e)
    for args in [(), (torch._C._autograd.DeviceType.CUDA, 0)]:
        try: print("has_multicast_support", args, sm._SymmetricMemory.has_multicast_support(*args))
        except Exception as e: print("has_mc err", args, repr(e)[:200])
    try: print("backend", sm.get_backend(torch.device("cuda",0)))
    except Exception as e: print("backend err", e)
dist.barrier(); dist.destroy_process_group()
