# Same import path as the reference (flash_attention/__init__.py:1):
#   from flash_attention import flash_attn_func
from flash_attention_cute_b200.flash_attention import flash_attn_func
from flash_attention_cute_b200.varlen import flash_attn_varlen_func
from flash_attention_cute_b200.paged import flash_attn_decode_paged

__all__ = ["flash_attn_func", "flash_attn_varlen_func", "flash_attn_decode_paged"]
