from flash_attention_cute_b200.models.patch_qwen2 import patch_attn  # noqa: F401
