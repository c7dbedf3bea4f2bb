"""B200 (sm_100a) FlashAttention-2 forward, drop-in for izmttk/flash_attention_cute's hot path.

``flash_attn_func`` is re-exported exactly as the reference's flash_attention/__init__.py:1 does.
"""
from .flash_attention import flash_attn_func
from .varlen import flash_attn_varlen_func
from .paged import flash_attn_decode_paged

__all__ = ["flash_attn_func", "flash_attn_varlen_func", "flash_attn_decode_paged"]
