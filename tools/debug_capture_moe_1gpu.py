"""Does a single-GPU MoE training step capture into a CUDA graph?  Stage by stage, each in its own try block.

    python tools/debug_capture_moe_1gpu.py [stage ...]      stages: route local_fwd local_fwdbwd step

Runs as its own process (a failed capture can leave the caching allocator in capture state and abort at interpreter
exit, which must not happen inside the test suite).  Prints one JSON line per stage.
"""
import json
import os
import sys
import traceback

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from scaletorch_b200.data import MicroBatchDataLoader  # noqa: E402
from scaletorch_b200.trainer import GraphedTrainStep, ScaleTorchArguments, create_model, create_optimizer  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)


def capture(fn, warm: int = 2):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(warm):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, capture_error_mode="thread_local"):
        out = fn()
    g.replay()
    torch.cuda.synchronize()
    return out


def main() -> None:
    stages = sys.argv[1:] or ["route", "local_fwd", "local_fwdbwd", "step"]
    torch.manual_seed(0)
    cfg = ScaleTorchArguments(model_name_or_path="tiny-qwen3-moe", micro_batch_size=2, sequence_length=256,
                              gradient_accumulation_steps=1, synthetic_data=True, dtype="bfloat16", learning_rate=1e-3,
                              max_grad_norm=1.0, seed=3)
    model, mc = create_model(cfg, torch.bfloat16, dev)
    model.train()
    from scaletorch_b200.models.model_qwen3_moe import MoELayer
    from scaletorch_b200.ops.moe_routing import route

    layer = next(m for m in model.modules() if isinstance(m, MoELayer))
    x = torch.randn(512, mc.hidden_size, device=dev).bfloat16().requires_grad_()
    gout = torch.randn(512, mc.hidden_size, device=dev).bfloat16()
    logits = torch.randn(512, mc.num_experts, device=dev).bfloat16().requires_grad_()

    def st_route():
        w, idx, pos, counts, aux = route(logits, mc.num_experts_per_tok, 0.001, True)
        return torch.autograd.grad(w.sum() + aux, logits)[0]

    def st_local_fwd():
        with torch.no_grad():
            return layer(x)

    def st_local_fwdbwd():
        y = layer(x)
        return torch.autograd.grad(y, x, gout)[0]

    table = {"route": st_route, "local_fwd": st_local_fwd, "local_fwdbwd": st_local_fwdbwd}
    for name in stages:
        if name == "step":
            continue
        try:
            capture(table[name])
            print(json.dumps({"stage": name, "captured": True}), flush=True)
        except Exception as e:
            print(json.dumps({"stage": name, "captured": False, "error": f"{type(e).__name__}: {str(e).splitlines()[0]}",
                              "where": traceback.format_exc().splitlines()[-6:]}), flush=True)
            return                                            # the allocator may be stuck in capture state: stop here
    if "step" in stages:
        opt = create_optimizer(model, cfg, dev)
        loader = MicroBatchDataLoader(2, 256, gradient_accumulation_steps=1, synthetic=True, vocab_size=mc.vocab_size,
                                      seed=5, pin_memory=True, num_samples=256)
        graphed = GraphedTrainStep(model, opt, cfg, dev, warmup=2)
        losses = [float(graphed(loader).item()) for _ in range(6)]
        print(json.dumps({"stage": "step", "captured": graphed.graph is not None, "losses": losses}), flush=True)


if __name__ == "__main__":
    main()
    sys.stdout.flush()
    os._exit(0)        # skip interpreter finalisation: see the module docstring
