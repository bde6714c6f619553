from .base import BaseAttention, validate_attention_inputs
from .gqa import GroupQueryAttention
from .mha import MultiHeadAttention
from .mla import MultiHeadLatentAttention
from .mqa import MultiQueryAttention

__all__ = ["BaseAttention", "GroupQueryAttention", "MultiHeadAttention", "MultiHeadLatentAttention", "MultiQueryAttention",
           "validate_attention_inputs"]
