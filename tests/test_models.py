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


@pytest.mark.gpu
def test_full_width_llama_layers_patched_vs_eager(cuda_device):
    """BASELINE config 5 at full width: two Llama-3-8B layers (hidden 4096, 32 q / 8 kv heads of 128,
    intermediate 14336), random-init bf16, S = 2048 — logits of the patched model against HF eager,
    and the "free view" after the op (reference models/rope_attn_fwd.py:62,117: O inherits q's
    [B,S,H,D]-memory strides, so transpose -> reshape -> contiguous copies nothing)."""
    from transformers import LlamaConfig, LlamaForCausalLM
    from transformers.models.llama.modeling_llama import LlamaAttention

    import flash_attention_cute_b200.models.rope_attn_fwd as raf
    from flash_attention_cute_b200 import _lib
    from models.patch_llama import patch_attn

    cfg = LlamaConfig(hidden_size=4096, num_hidden_layers=2, num_attention_heads=32, num_key_value_heads=8,
                      intermediate_size=14336, vocab_size=4096, rope_theta=500000.0, max_position_embeddings=4096,
                      attn_implementation="eager")
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg).to(device=cuda_device, dtype=torch.bfloat16).eval()
    ids = torch.randint(0, 4096, (2, 2048), generator=torch.Generator().manual_seed(2)).to(cuda_device)
    with torch.no_grad():
        eager = model(ids).logits.float()
        model32 = LlamaForCausalLM(cfg).to(device=cuda_device, dtype=torch.float32).eval()
        model32.load_state_dict({k: v.float() for k, v in model.state_dict().items()})
        truth = model32(ids).logits        # same (bf16-valued) weights, fp32 arithmetic
        del model32
    original = LlamaAttention.forward
    real_op = raf.flash_attn_func
    views = []

    def spying_op(q, k, v, softmax_scale=None, causal=False):
        o = real_op(q, k, v, softmax_scale=softmax_scale, causal=causal)
        flat = o.transpose(1, 2).reshape(o.size(0), o.size(2), -1)
        views.append(o.stride() == q.stride() and flat.is_contiguous() and flat.contiguous().data_ptr() == o.data_ptr())
        return o

    raf.flash_attn_func = spying_op
    try:
        patch_attn()
        n0 = _lib.launch_count()
        with torch.no_grad():
            got = model(ids).logits.float()
        _lib.check_device(0)
        assert _lib.launch_count() - n0 == 4  # per layer: fused RoPE + attention
    finally:
        LlamaAttention.forward = original
        raf.flash_attn_func = real_op
    assert views == [True, True], "the reshape after the op must stay a free view"
    # both bf16 models against the fp32 evaluation of the same weights: the patched model must be
    # no further from the truth than HF's own eager bf16 path (plus a quarter)
    err_ours = ((got - truth).norm() / truth.norm()).item()
    err_eager = ((eager - truth).norm() / truth.norm()).item()
    assert err_ours <= 1.25 * err_eager + 1e-3, (err_ours, err_eager)
