from flash_attention_cute_b200.models.rope_attn_fwd import (  # noqa: F401
    _flash_attention_forward,
    apply_rotary_pos_emb,
    attention_forward,
    rotate_half,
)
