"""Host-side cost of one operator call: wall time per call of a problem so small
that the GPU is never the limit, for the public op (`flash_attn_func` -> torch custom op -> shim ->
ctypes -> C ABI), for the shim alone and for the bare C ABI call (`fa_sm100_fwd` with a filled
params struct: the launch-plan cache hit path, no tensor-map encode)."""
import ctypes
import json
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from flash_attention import flash_attn_func  # noqa: E402
from flash_attention_cute_b200 import _lib  # noqa: E402
from flash_attention_cute_b200.load_cpp_extention import fill_params, load_extension  # noqa: E402


def per_call_us(fn, n=3000):
    for _ in range(200):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    return (t1 - t0) / n * 1e6


def main():
    dev = torch.device("cuda:0")
    ext = load_extension()
    lib = _lib.load()
    out = {}
    for name, (B, Hq, Hkv, Sq, Sk, D) in {"prefill 1x8x256x128": (1, 8, 8, 256, 256, 128),
                                          "decode 16x64/8x1024x128": (16, 64, 8, 1, 1024, 128)}.items():
        q = torch.randn(B, Hq, Sq, D, device=dev).bfloat16()
        k = torch.randn(B, Hkv, Sk, D, device=dev).bfloat16()
        v = torch.randn(B, Hkv, Sk, D, device=dev).bfloat16()
        o = torch.empty_like(q)
        rec = {"public_op_us": per_call_us(lambda: flash_attn_func(q, k, v, causal=True)),
               "shim_us": per_call_us(lambda: ext.flash_attention_fwd(q, k, v, D ** -0.5, True))}
        if Sq > 1:
            p = fill_params(q, k, v, o, D ** -0.5, True)
            stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            ref = ctypes.byref(p)
            rec["c_abi_us"] = per_call_us(lambda: lib.fa_sm100_fwd(ref, stream))
        out[name] = rec
    _lib.check_device(0)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
