"""Replacement ``forward`` for HF Llama / Qwen2 attention modules.

Same entry points and control flow as reference models/rope_attn_fwd.py:
``rotate_half`` (:8-12), ``apply_rotary_pos_emb`` (:14-38), ``_flash_attention_forward`` (:40-64)
and ``attention_forward`` (:66-120): q/k/v projections viewed ``[B,S,H,D] -> transpose(1,2)``
(strided views, never copied), RoPE, optional KV-cache update, one ``flash_attn_func`` call with
``causal=module.is_causal`` and ``softmax_scale=module.scaling``, ``transpose(1,2)``, reshape,
``o_proj``.  Because the op's output inherits q's strides, the trailing
``reshape(B,S,-1).contiguous()`` stays a free view.

Two reference quirks are made robust for the transformers version in this image (5.x), see
SURVEY.md §8(b): the cache kwarg may arrive as ``past_key_values`` (plural), and
``config.use_sliding_window`` does not exist on ``LlamaConfig``.  As in the reference the
attention mask, dropout, sliding window and softcap are accepted and ignored (:45-61).
"""
from typing import Optional, Tuple

import torch
from torch import nn

from flash_attention_cute_b200.flash_attention import flash_attn_func


def rotate_half(x):
    """Rotates half the hidden dims of the input."""
    half = x.shape[-1] // 2
    return torch.cat((-x[..., half:], x[..., :half]), dim=-1)


def _fused_rope_applies(q, k, cos, sin, unsqueeze_dim) -> bool:
    return (q.is_cuda and unsqueeze_dim == 1 and q.dim() == 4 and k.dim() == 4 and cos.dim() == 3
            and q.dtype in (torch.float16, torch.bfloat16)
            and q.dtype == k.dtype == cos.dtype == sin.dtype
            and q.size(-1) % 16 == 0 and cos.size(-1) == q.size(-1) and cos.shape == sin.shape
            and cos.size(1) == q.size(2) == k.size(2) and cos.size(0) in (1, q.size(0))
            and not (q.requires_grad or k.requires_grad))


def apply_rotary_pos_emb(q, k, cos, sin, position_ids=None, unsqueeze_dim=1):
    """q_embed = q·cos + rotate_half(q)·sin (same for k); cos/sin are [B, S, D].

    On the GPU, for the layout the model path produces (16-bit, [B,H,S,D] views, cos/sin of the
    same dtype), this is ONE kernel (`fa_sm100_rope`, csrc/fa_sm100_rope.cuh) that is bit-equal to
    the expression below; the six elementwise passes and five temporaries of the eager form go
    away.  Anything else (CPU tensors of the host-logic tests, autograd, exotic
    broadcasting) evaluates the reference's expression as written."""
    if _fused_rope_applies(q, k, cos, sin, unsqueeze_dim):
        from flash_attention_cute_b200.load_cpp_extention import load_extension

        return load_extension().rope(q, k, cos, sin)
    cos = cos.unsqueeze(unsqueeze_dim)
    sin = sin.unsqueeze(unsqueeze_dim)
    return (q * cos) + (rotate_half(q) * sin), (k * cos) + (rotate_half(k) * sin)


def _flash_attention_forward(module: nn.Module, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                             attention_mask: Optional[torch.Tensor], dropout: float = 0.0,
                             scaling: Optional[float] = None, sliding_window: Optional[int] = None,
                             softcap: Optional[float] = None, **kwargs) -> Tuple[torch.Tensor, None]:
    kwargs.pop("is_causal", None)
    attn_output = flash_attn_func(query, key, value, causal=module.is_causal, softmax_scale=scaling)
    return attn_output.transpose(1, 2), None


def attention_forward(self: nn.Module, hidden_states: torch.Tensor,
                      position_embeddings: Tuple[torch.Tensor, torch.Tensor],
                      attention_mask: Optional[torch.Tensor] = None,
                      past_key_value=None, cache_position: Optional[torch.LongTensor] = None, **kwargs):
    if past_key_value is None:
        past_key_value = kwargs.pop("past_key_values", None)
    else:
        kwargs.pop("past_key_values", None)

    input_shape = hidden_states.shape[:-1]
    hidden_shape = (*input_shape, -1, self.head_dim)
    # [B, H, S, D] views over [B, S, H, D] memory
    query_states = self.q_proj(hidden_states).view(hidden_shape).transpose(1, 2)
    key_states = self.k_proj(hidden_states).view(hidden_shape).transpose(1, 2)
    value_states = self.v_proj(hidden_states).view(hidden_shape).transpose(1, 2)

    cos, sin = position_embeddings
    query_states, key_states = apply_rotary_pos_emb(query_states, key_states, cos, sin)

    if past_key_value is not None:
        cache_kwargs = {"sin": sin, "cos": cos, "cache_position": cache_position}
        key_states, value_states = past_key_value.update(key_states, value_states, self.layer_idx, cache_kwargs)

    sliding_window = None
    if (getattr(self.config, "use_sliding_window", False)
            and getattr(self.config, "sliding_window", None) is not None
            and self.layer_idx >= getattr(self.config, "max_window_layers", 0)):
        sliding_window = self.config.sliding_window  # read, never honoured — as in the reference

    attn_output, attn_weights = _flash_attention_forward(
        self, query_states, key_states, value_states, attention_mask,
        dropout=0.0 if not self.training else getattr(self, "attention_dropout", 0.0),
        scaling=self.scaling, sliding_window=sliding_window, **kwargs)

    attn_output = attn_output.reshape(*input_shape, -1).contiguous()
    attn_output = self.o_proj(attn_output)
    return attn_output, attn_weights
