"""Python entry point and torch custom-op registration.

Mirrors reference flash_attention/flash_attention.py:
  * ``flash_attn_func(q, k, v, softmax_scale=None, causal=False)``           (:46-53)
  * custom op ``flash_attention::forward`` with schema
    ``(Tensor q, Tensor k, Tensor v, float softmax_scale=None, bool causal=False) -> Tensor``,
    ``mutates_args=()``                                                        (:6-7)
  * ``"cuda"`` kernel: pad D to a multiple of 8, force unit last-dim stride, call the
    extension, slice the padding off                                           (:17-38)
  * fake/meta kernel ``empty_like(q)``                                         (:40-43)

One deliberate difference (BASELINE.json north_star: "no CPU fallback"): the reference's default
implementation prints a notice and calls ``F.scaled_dot_product_attention`` (:13-15); here a
non-CUDA call raises ``NotImplementedError``.
"""
import torch

from .load_cpp_extention import load_extension

flash_attention_cuda = load_extension()


@torch.library.custom_op("flash_attention::forward", mutates_args=())
def flash_attention_forward(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor,
                            softmax_scale: float = None, causal: bool = False) -> torch.Tensor:
    # q: [batch_size, n_heads, q_seq_len,  d]
    # k: [batch_size, n_heads, kv_seq_len, d]
    # v: [batch_size, n_heads, kv_seq_len, d]
    raise NotImplementedError(
        "flash_attention::forward is implemented for CUDA (sm_100a / B200) tensors only; "
        f"got device {q.device}. There is no CPU or PyTorch fallback."
    )


@torch.library.register_kernel("flash_attention::forward", "cuda")
def flash_attention_forward_cuda(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor,
                                 softmax_scale: float = None, causal: bool = False) -> torch.Tensor:
    head_dim = q.size(3)
    if softmax_scale is None:
        softmax_scale = head_dim ** -0.5

    need_padding = False
    if head_dim % 8 != 0:
        need_padding = True
        pad = [0, 8 - head_dim % 8]
        q = torch.nn.functional.pad(q, pad)
        k = torch.nn.functional.pad(k, pad)
        v = torch.nn.functional.pad(v, pad)
    q = q.contiguous() if q.stride(3) != 1 else q
    k = k.contiguous() if k.stride(3) != 1 else k
    v = v.contiguous() if v.stride(3) != 1 else v
    attn = flash_attention_cuda.flash_attention_fwd(q, k, v, softmax_scale, causal)
    if need_padding:
        attn = attn[:, :, :, :head_dim]
    return attn


@torch.library.register_fake("flash_attention::forward")
def flash_attention_forward_fake(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor,
                                 softmax_scale: float = None, causal: bool = False) -> torch.Tensor:
    return torch.empty_like(q)


def flash_attn_func(q, k, v, softmax_scale=None, causal=False):
    # q: [batch_size, n_heads, q_seq_len,  d]
    # k: [batch_size, n_heads, kv_seq_len, d]
    # v: [batch_size, n_heads, kv_seq_len, d]
    softmax_scale = (q.size(-1) ** -0.5) if softmax_scale is None else softmax_scale
    return torch.ops.flash_attention.forward(q, k, v, softmax_scale, causal)
