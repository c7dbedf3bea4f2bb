from flash_attention_cute_b200.models.patch_llama import patch_attn  # noqa: F401
