"""Whole-step CUDA-graph replay must reproduce the eager training trajectory (same kernels, same order)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(use_graph: bool, steps: int = 6):
    from scaletorch_b200.data import MicroBatchDataLoader
    from scaletorch_b200.trainer import GraphedTrainStep, ScaleTorchArguments, create_model, create_optimizer, training_step

    torch.manual_seed(0)
    cfg = ScaleTorchArguments(model_name_or_path="tiny-qwen3", micro_batch_size=2, sequence_length=256,
                              gradient_accumulation_steps=2, synthetic_data=True, dtype="bfloat16", learning_rate=1e-3,
                              max_grad_norm=1.0, seed=3)
    device = torch.device("cuda", 0)
    model, mc = create_model(cfg, torch.bfloat16, device)
    model.train()
    opt = create_optimizer(model, cfg, device)
    loader = MicroBatchDataLoader(2, 256, gradient_accumulation_steps=2, synthetic=True, vocab_size=mc.vocab_size, seed=5,
                                  pin_memory=True, num_samples=512)
    losses = []
    graphed = GraphedTrainStep(model, opt, cfg, device, warmup=2) if use_graph else None
    for i in range(steps):
        for g in opt.param_groups:
            g["lr"] = 1e-3 * (1 + i)             # a changing LR must reach the replayed optimiser kernel
        if graphed is not None:
            loss = graphed(loader)
        else:
            opt.zero_grad()
            loss = training_step(model, loader, device, torch.bfloat16, cfg)
            opt.step()
        losses.append(float(loss.item()))
    assert graphed is None or graphed.graph is not None
    return losses, opt.step_count


def test_graphed_step_matches_eager():
    eager, n0 = _run(False)
    graph, n1 = _run(True)
    assert n0 == n1 == 6
    for a, b in zip(eager, graph):
        assert abs(a - b) < 2e-3 * max(1.0, abs(a)), (eager, graph)
    assert eager[-1] < eager[0]
