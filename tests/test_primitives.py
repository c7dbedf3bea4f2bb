"""GPU: the UMMA / TMA / TMEM descriptor recipes used by the forward kernel, in isolation
(fa_sm100_selftest_umma): S = Q·Kᵀ (SS mode, K-major SW128) and O = T(S·c)·V (TS mode, A from
TMEM, V MN-major SW128)."""
import ctypes

import pytest
import torch

from flash_attention_cute_b200 import _lib

pytestmark = pytest.mark.gpu


def run_selftest(dev, D, dtype, seed=0, p_scale=0.125, v_lbo=0, v_sbo=0):
    g = torch.Generator(device=dev).manual_seed(seed)
    q = torch.randn(128, D, generator=g, device=dev).to(dtype)
    k = torch.randn(128, D, generator=g, device=dev).to(dtype)
    v = torch.randn(128, D, generator=g, device=dev).to(dtype)
    s = torch.full((128, 128), float("nan"), device=dev)
    o = torch.full((128, D), float("nan"), device=dev)
    lib = _lib.load()
    rc = lib.fa_sm100_selftest_umma(q.data_ptr(), k.data_ptr(), v.data_ptr(), s.data_ptr(), o.data_ptr(),
                                    D, _lib.FA_SM100_BF16 if dtype == torch.bfloat16 else _lib.FA_SM100_FP16,
                                    p_scale, v_lbo, v_sbo, dev.index or 0,
                                    ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    _lib.raise_for_status(rc)
    _lib.check_device(dev.index or 0)
    s_ref = q.float() @ k.float().T
    p_ref = (s * p_scale).to(dtype).float()  # P is rounded from the kernel's own S
    o_ref = p_ref @ v.float()
    return s, s_ref, o, o_ref


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("D", [128, 64, 96, 32])
def test_umma_qk_and_pv(cuda_device, D, dtype):
    s, s_ref, o, o_ref = run_selftest(cuda_device, D, dtype)
    torch.testing.assert_close(s, s_ref, rtol=1e-4, atol=1e-3)
    torch.testing.assert_close(o, o_ref, rtol=1e-4, atol=2e-3)
