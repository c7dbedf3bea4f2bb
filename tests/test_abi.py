"""CPU: the C-ABI library loads, exports every symbol include/fa_sm100.h declares, validates
arguments like the reference host function, and plans launches (no compute without a GPU)."""
import ctypes
import re
from pathlib import Path

import pytest
import torch

from flash_attention_cute_b200 import _build, _lib
from flash_attention_cute_b200.load_cpp_extention import fill_params, load_extension

ROOT = Path(__file__).resolve().parent.parent


def test_library_builds_and_loads():
    lib = _lib.load()
    assert _build.LIB.exists()
    assert b"sm_100a" in lib.fa_sm100_build_info()


def test_exports_every_declared_symbol():
    header = (ROOT / "include" / "fa_sm100.h").read_text()
    declared = set(re.findall(r"\b(fa_sm100_[a-z_0-9]+)\s*\(", header))
    declared -= {"fa_sm100_fwd_params", "fa_sm100_varlen_params", "fa_sm100_paged_decode_params", "fa_sm100_rope_params"}
    assert declared == set(_lib.EXPORTED_SYMBOLS)
    lib = _lib.load()
    for name in declared:
        assert getattr(lib, name) is not None


def test_struct_layout_matches_header():
    # 4 ptr + 6 i64 + 12 i64 + float + 3 i32 + u32 = 32 + 48 + 96 + 20 = 196 -> padded to 200
    assert ctypes.sizeof(_lib.FwdParams) == 200
    assert _lib.FwdParams.q_strides.offset == 80
    assert _lib.FwdParams.softmax_scale.offset == 176


def _plan(q, k, causal=False, flags=0):
    return load_extension().plan(q, k, k, causal, flags)


def test_plan_configs():
    bf = torch.bfloat16
    # C2: B=4,H=32,S=4096,D=128
    q = torch.empty(4, 32, 4096, 128, dtype=bf)
    p = _plan(q, q)
    assert p == dict(grid=4 * 32 * 16, smem_bytes=p["smem_bytes"], pack_g=1, rows_per_head=128)
    assert p["smem_bytes"] <= 232448
    # C3: GQA 32/8, S=8192, forced packing -> 4 heads x 32 positions per M tile
    q = torch.empty(1, 32, 8192, 128, dtype=bf)
    k = torch.empty(1, 8, 8192, 128, dtype=bf)
    p = _plan(q, k, True, _lib.FLAG_FORCE_PACK_GQA)
    assert (p["pack_g"], p["rows_per_head"], p["grid"]) == (4, 32, 8 * 128)
    # C4: Qwen2-0.5B 14/2 -> 7 heads x 18 positions (126 of 128 rows used)
    q = torch.empty(1, 14, 512, 64, dtype=torch.float16)
    k = torch.empty(1, 2, 512, 64, dtype=torch.float16)
    p = _plan(q, k, True, _lib.FLAG_FORCE_PACK_GQA)
    assert (p["pack_g"], p["rows_per_head"]) == (7, 18)
    # decode: packed by default
    p = _plan(q[:, :, :1], k, True)
    assert (p["pack_g"], p["grid"]) == (7, 2)


@pytest.mark.parametrize("mutate,msg", [
    (lambda p: setattr(p, "heads_kv", 3), "multiple of number of heads"),
    (lambda p: setattr(p, "headdim", 136), "hidden dimension <= 128"),
    (lambda p: setattr(p, "headdim", 60), "multiple of 8"),
    (lambda p: setattr(p, "dtype", 7), "fp16 or bf16"),
    (lambda p: setattr(p, "seqlen_q", 0), "at least one element"),
    (lambda p: setattr(p, "q", 8), "16-byte aligned"),
])
def test_validation_errors(mutate, msg):
    q = torch.empty(1, 4, 64, 64, dtype=torch.float16)
    p = fill_params(q, q, q, q, 1.0, False)
    mutate(p)
    lib = _lib.load()
    rc = lib.fa_sm100_plan(ctypes.byref(p), None, None, None, None)
    assert rc == 1
    assert msg in _lib.last_error()


def test_op_schema_and_fake_and_no_cpu_fallback():
    from flash_attention import flash_attn_func

    schema = str(torch.ops.flash_attention.forward.default._schema)
    assert schema == ("flash_attention::forward(Tensor q, Tensor k, Tensor v, "
                      "float softmax_scale=None, bool causal=False) -> Tensor")
    q = torch.randn(1, 2, 8, 16)
    with pytest.raises(NotImplementedError):
        flash_attn_func(q, q, q)
    # fake impl: empty_like(q) keeps q's (transposed) strides
    from torch._subclasses.fake_tensor import FakeTensorMode

    with FakeTensorMode():
        qf = torch.empty(2, 16, 4, 8, dtype=torch.float16).transpose(1, 2)
        out = torch.ops.flash_attention.forward(qf, qf, qf, 1.0, True)
        assert out.shape == qf.shape and out.stride() == qf.stride()


def test_product_never_imports_oracle():
    for f in (ROOT / "flash_attention_cute_b200").rglob("*.py"):
        assert "oracle" not in f.read_text(), f
    for f in list((ROOT / "flash_attention").glob("*.py")) + list((ROOT / "models").glob("*.py")):
        assert "oracle" not in f.read_text(), f
