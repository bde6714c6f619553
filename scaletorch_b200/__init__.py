"""scaletorch_b200 -- a Blackwell (B200 / sm_100a) native 5-D parallel LLM training framework with the
capabilities of jianzhnie/ScaleTorch: ``ProcessGroupManager`` over the ``[DP, PP, CP, EP, TP]`` grid,
``HfArgumentParser`` training dataclasses, Llama / Qwen3 / Qwen3-MoE, HF-compatible checkpoints --
on hand-written tcgen05 / TMA kernels and fused compute+collective kernels over NVLink peer memory.
"""
__version__ = "0.2.0"
