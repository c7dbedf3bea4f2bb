"""Reference models/patch_llama.py:4-5 — class-level monkey patch of LlamaAttention.forward."""
from transformers.models.llama.modeling_llama import LlamaAttention

from .rope_attn_fwd import attention_forward


def patch_attn():
    LlamaAttention.forward = attention_forward
