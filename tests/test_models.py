"""models/patch_llama.py and models/patch_qwen2.py (reference models/patch_*.py:4-5,
models/rope_attn_fwd.py:66-120).  CPU: host logic of the patched forward with the attention op
replaced by the oracle (test-only substitution).  GPU: tiny random-init models, patched vs HF eager."""
import pytest
import torch

from oracle.eager_attention import eager_attention


def tiny_config(kind):
    if kind == "llama":
        from transformers import LlamaConfig
        return LlamaConfig(hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
                           num_key_value_heads=2, vocab_size=512, max_position_embeddings=1024,
                           attn_implementation="eager")
    from transformers import Qwen2Config
    return Qwen2Config(hidden_size=448, intermediate_size=512, num_hidden_layers=2, num_attention_heads=7,
                       num_key_value_heads=1, vocab_size=512, max_position_embeddings=1024,
                       attn_implementation="eager")


def build(kind, dtype, device):
    from transformers import LlamaForCausalLM, Qwen2ForCausalLM
    torch.manual_seed(0)
    cls = LlamaForCausalLM if kind == "llama" else Qwen2ForCausalLM
    return cls(tiny_config(kind)).to(device=device, dtype=dtype).eval()


def attn_class(kind):
    if kind == "llama":
        from transformers.models.llama.modeling_llama import LlamaAttention
        return LlamaAttention
    from transformers.models.qwen2.modeling_qwen2 import Qwen2Attention
    return Qwen2Attention


def patcher(kind):
    if kind == "llama":
        from models.patch_llama import patch_attn
    else:
        from models.patch_qwen2 import patch_attn
    return patch_attn


@pytest.mark.parametrize("kind", ["llama", "qwen2"])
def test_patch_host_logic_cpu(kind, monkeypatch):
    import flash_attention_cute_b200.models.rope_attn_fwd as raf

    cls = attn_class(kind)
    original = cls.forward
    model = build(kind, torch.float32, "cpu")
    ids = torch.randint(0, 512, (2, 48), generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        want = model(ids).logits
    seen = {}

    def fake_op(q, k, v, softmax_scale=None, causal=False):
        seen["strides"] = (q.stride(), k.stride(), v.stride())
        seen["causal"] = causal
        out = torch.empty_like(q)  # the op's contract: O has q's strides
        out.copy_(eager_attention(q, k, v, softmax_scale, causal))
        return out

    monkeypatch.setattr(raf, "flash_attn_func", fake_op)
    try:
        patcher(kind)()
        assert cls.forward is raf.attention_forward
        with torch.no_grad():
            got = model(ids).logits
    finally:
        cls.forward = original
    torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-4)
    cfg = tiny_config(kind)
    D, Hq = cfg.hidden_size // cfg.num_attention_heads, cfg.num_attention_heads
    assert seen["causal"] is True
    assert seen["strides"][0] == (48 * Hq * D, D, Hq * D, 1)  # [B,S,H,D] memory seen as [B,H,S,D]


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["llama", "qwen2"])
def test_patched_model_matches_eager_gpu(cuda_device, kind):
    from flash_attention_cute_b200 import _lib

    cls = attn_class(kind)
    original = cls.forward
    model = build(kind, torch.bfloat16, cuda_device)
    ids = torch.randint(0, 512, (2, 300), generator=torch.Generator().manual_seed(1)).to(cuda_device)
    with torch.no_grad():
        want = model(ids).logits.float()
    n0 = _lib.launch_count()
    try:
        patcher(kind)()
        with torch.no_grad():
            got = model(ids).logits.float()
    finally:
        cls.forward = original
    _lib.check_device(0)
    assert _lib.launch_count() - n0 == 4  # per layer: one fused-RoPE launch + one attention launch
    rel = ((got - want).norm() / want.norm()).item()
    assert rel < 2e-2, rel
