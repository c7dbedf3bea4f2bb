"""Fused RoPE kernel (reference models/rope_attn_fwd.py:8-12, 14-38, 87-88).

The parity bar here is bit-exactness: the kernel reproduces PyTorch's 16-bit evaluation of
``q*cos + rotate_half(q)*sin`` rounding for rounding.  CPU: a numpy restatement of those roundings
against the reference expression (pins the arithmetic the kernel implements) and the dispatch
logic.  GPU: the kernel against the reference expression evaluated by torch on the same device."""
import numpy as np
import pytest
import torch


def ref_rotate_half(x):  # models/rope_attn_fwd.py:8-12
    x1 = x[..., : x.shape[-1] // 2]
    x2 = x[..., x.shape[-1] // 2:]
    return torch.cat((-x2, x1), dim=-1)


def ref_apply(q, k, cos, sin, unsqueeze_dim=1):  # models/rope_attn_fwd.py:34-38
    cos = cos.unsqueeze(unsqueeze_dim)
    sin = sin.unsqueeze(unsqueeze_dim)
    return (q * cos) + (ref_rotate_half(q) * sin), (k * cos) + (ref_rotate_half(k) * sin)


def make(B, Hq, Hkv, S, D, dtype, device, seed=0, bcast=False):
    g = torch.Generator(device="cpu").manual_seed(seed)
    # [B, S, H, D] memory viewed as [B, H, S, D] — what q_proj(...).view(...).transpose(1, 2) gives
    q = torch.randn(B, S, Hq, D, generator=g).to(dtype).to(device).transpose(1, 2)
    k = torch.randn(B, S, Hkv, D, generator=g).to(dtype).to(device).transpose(1, 2)
    pos = torch.arange(S, dtype=torch.float32)[None, :, None] + torch.arange(1 if bcast else B)[:, None, None] * 3.0
    inv = 1.0 / (10000.0 ** (torch.arange(0, D, 2, dtype=torch.float32) / D))
    ang = torch.cat([pos * inv, pos * inv], dim=-1)
    return q, k, ang.cos().to(dtype).to(device), ang.sin().to(dtype).to(device)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16], ids=["fp16", "bf16"])
def test_three_roundings_restate_the_reference(dtype):
    """out = rn(rn(x·cos) + rn(x_rot·sin)) evaluated in fp32 with explicit roundings equals the
    reference expression evaluated by torch in `dtype` (CPU)."""
    q, k, cos, sin = make(2, 4, 2, 33, 64, dtype, "cpu", seed=3)
    want, _ = ref_apply(q, k, cos, sin)
    rn = lambda t: t.to(dtype).float()
    c, s = cos.unsqueeze(1).float(), sin.unsqueeze(1).float()
    x = q.float()
    half = x.shape[-1] // 2
    xr = torch.cat((-x[..., half:], x[..., :half]), dim=-1)
    got = (rn(x * c) + rn(xr * s)).to(dtype)
    assert torch.equal(got, want)


def test_model_path_keeps_reference_expression_on_cpu():
    import flash_attention_cute_b200.models.rope_attn_fwd as raf

    q, k, cos, sin = make(1, 4, 2, 16, 32, torch.bfloat16, "cpu")
    a = raf.apply_rotary_pos_emb(q, k, cos, sin)
    b = ref_apply(q, k, cos, sin)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    assert not raf._fused_rope_applies(q, k, cos, sin, 1)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16], ids=["fp16", "bf16"])
@pytest.mark.parametrize("B,Hq,Hkv,S,D,bcast", [
    (2, 32, 8, 512, 128, False),
    (1, 14, 2, 1000, 64, True),
    (3, 8, 8, 77, 128, True),
    (2, 4, 1, 1, 64, False),      # decode step
    (1, 2, 2, 130, 256, False),
])
def test_bit_exact_against_reference_expression(cuda_device, dtype, B, Hq, Hkv, S, D, bcast):
    from flash_attention_cute_b200 import _lib
    from flash_attention_cute_b200.load_cpp_extention import load_extension

    q, k, cos, sin = make(B, Hq, Hkv, S, D, dtype, cuda_device, seed=7, bcast=bcast)
    before = _lib.launch_count()
    qo, ko = load_extension().rope(q, k, cos, sin)
    _lib.check_device(q.device.index or 0)
    assert _lib.launch_count() == before + 1
    rq, rk = ref_apply(q, k, cos, sin)
    assert qo.stride() == q.stride() and ko.stride() == k.stride()
    assert torch.equal(qo, rq) and torch.equal(ko, rk)


@pytest.mark.gpu
def test_model_path_uses_the_kernel_and_matches(cuda_device):
    import flash_attention_cute_b200.models.rope_attn_fwd as raf
    from flash_attention_cute_b200 import _lib

    q, k, cos, sin = make(2, 8, 2, 256, 128, torch.bfloat16, cuda_device, seed=1)
    assert raf._fused_rope_applies(q, k, cos, sin, 1)
    before = _lib.launch_count()
    a = raf.apply_rotary_pos_emb(q, k, cos, sin)
    assert _lib.launch_count() == before + 1
    b = ref_apply(q, k, cos, sin)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    # contiguous inputs and fp32 tensors take the expression path / still agree
    a32 = raf.apply_rotary_pos_emb(q.float(), k.float(), cos.float(), sin.float())
    b32 = ref_apply(q.float(), k.float(), cos.float(), sin.float())
    assert torch.equal(a32[0], b32[0])
