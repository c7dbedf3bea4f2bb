"""First-contact diagnostics on a B200: primitive self-test with descriptor variants, then a few
forward cases with error numbers.  Writes gpurun_out/diag.json.  Not part of the product."""
import json
import sys
import time
import traceback
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

out = {"device": torch.cuda.get_device_name(0), "results": []}


def rec(name, **kw):
    kw["name"] = name
    out["results"].append(kw)
    print(json.dumps(kw), flush=True)


def main():
    from test_primitives import run_selftest
    from common import errors, make_qkv_fast
    from oracle.eager_attention import eager_attention, eager_attention_chunked, attention_flops
    from flash_attention_cute_b200 import _lib
    from flash_attention_cute_b200.load_cpp_extention import load_extension

    dev = torch.device("cuda:0")
    for D in (128, 64):
        for dtype in (torch.bfloat16, torch.float16):
            for (lbo, sbo) in ((0, 0),):
                try:
                    s, s_ref, o, o_ref = run_selftest(dev, D, dtype, v_lbo=lbo, v_sbo=sbo)
                    rec("selftest", D=D, dtype=str(dtype), lbo=lbo, sbo=sbo,
                        s_err=(s - s_ref).abs().max().item(), o_err=(o - o_ref).abs().max().item(),
                        s_nan=int(torch.isnan(s).sum()), o_nan=int(torch.isnan(o).sum()))
                except Exception as e:
                    rec("selftest", D=D, dtype=str(dtype), lbo=lbo, sbo=sbo, error=repr(e))
    ext = load_extension()
    cases = [
        (1, 1, 1, 128, 128, 128, False, 0), (1, 1, 1, 256, 256, 128, False, 0), (1, 2, 2, 512, 512, 128, False, 0),
        (1, 2, 2, 512, 512, 128, True, 0), (1, 4, 4, 1000, 1000, 64, True, 0), (2, 8, 2, 1024, 1024, 128, True, 1),
        (2, 8, 2, 1024, 1024, 128, True, 2), (1, 14, 2, 640, 640, 64, True, 2), (1, 14, 2, 1, 512, 64, True, 0),
        (1, 4, 4, 300, 200, 64, True, 0), (4, 32, 32, 4096, 4096, 128, False, 0), (1, 32, 8, 8192, 8192, 128, True, 1),
        (1, 32, 8, 8192, 8192, 128, True, 2),
    ]
    for (B, Hq, Hkv, Sq, Sk, D, causal, flags) in cases:
        for dtype in (torch.bfloat16,):
            try:
                q, k, v = make_qkv_fast(B, Hq, Hkv, Sq, Sk, D, dtype, 1, dev)
                o = ext.flash_attention_fwd(q, k, v, D ** -0.5, causal, flags)
                _lib.check_device(0)
                nb = min(B, 1); nh = min(Hkv, 2); G = Hq // Hkv
                ref = eager_attention_chunked(q[:nb, :nh * G], k[:nb, :nh], v[:nb, :nh], None, causal, out_dtype=torch.float32)
                ma, rl = errors(o[:nb, :nh * G], ref)
                # timing
                for _ in range(3):
                    ext.flash_attention_fwd(q, k, v, D ** -0.5, causal, flags)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
                n = 10
                e0.record()
                for _ in range(n):
                    ext.flash_attention_fwd(q, k, v, D ** -0.5, causal, flags)
                e1.record(); torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / n
                tf = attention_flops(B, Hq, Sq, Sk, D, causal) / ms / 1e9
                rec("fwd", shape=[B, Hq, Hkv, Sq, Sk, D], causal=causal, flags=flags, max_abs=ma, rel_l2=rl,
                    nan=int(torch.isnan(o.float()).sum()), ms=ms, tflops=tf)
            except Exception as e:
                rec("fwd", shape=[B, Hq, Hkv, Sq, Sk, D], causal=causal, flags=flags, error=repr(e),
                    tb=traceback.format_exc()[-600:])
                if "CUDA" in repr(e) or "ABORT" in repr(e):
                    break


if __name__ == "__main__":
    try:
        main()
    finally:
        Path(ROOT / "gpurun_out").mkdir(exist_ok=True)
        (ROOT / "gpurun_out" / "diag.json").write_text(json.dumps(out, indent=1))
