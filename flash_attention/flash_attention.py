# Same module path as the reference (flash_attention/flash_attention.py).
from flash_attention_cute_b200.flash_attention import (  # noqa: F401
    flash_attention_cuda,
    flash_attention_forward,
    flash_attention_forward_cuda,
    flash_attention_forward_fake,
    flash_attn_func,
)
