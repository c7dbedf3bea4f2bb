"""Reference models/patch_qwen2.py:4-5 — class-level monkey patch of Qwen2Attention.forward."""
from transformers.models.qwen2.modeling_qwen2 import Qwen2Attention

from .rope_attn_fwd import attention_forward


def patch_attn():
    Qwen2Attention.forward = attention_forward
