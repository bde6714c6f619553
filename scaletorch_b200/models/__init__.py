from .attention_utils import (RMSNorm, apply_rotary_pos_emb, get_attention_backend, get_cos_sin,
                              register_attention_backend)
from .llama import Llama
from .model_qwen3 import Qwen3
from .model_qwen3_moe import Qwen3MoE

__all__ = ["Llama", "Qwen3", "Qwen3MoE", "RMSNorm", "apply_rotary_pos_emb", "get_attention_backend", "get_cos_sin",
           "register_attention_backend"]
