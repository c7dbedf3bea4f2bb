"""Diagnosis helper: runs one named long-sequence case per process (a device fault kills the context)."""
import subprocess
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

CASES = {
    "fixed_bhsd_32k": "fixed", "fixed_bshd_32k": "fixed_view", "fixed_bshd_16k": "fixed_view16",
    "varlen_32k": [32768], "varlen_32k_1k": [32768, 1024], "varlen_1k_32k": [1024, 32768],
    "varlen_33x1k": [1024] * 33, "varlen_16k_16k_1k": [16384, 16385, 1024], "varlen_full": [32768] + [1024] * 32,
}


def run(name):
    import torch
    from flash_attention import flash_attn_func, flash_attn_varlen_func
    from flash_attention_cute_b200 import _lib

    spec = CASES[name]
    hq, hkv, d = 32, 8, 128
    g = torch.Generator(device="cuda").manual_seed(1)
    if isinstance(spec, str):
        s = 16384 if spec.endswith("16") else 32768
        if spec == "fixed":
            q = torch.randn(1, hq, s, d, generator=g, device="cuda").bfloat16()
            k = torch.randn(1, hkv, s, d, generator=g, device="cuda").bfloat16()
        else:
            q = torch.randn(1, s, hq, d, generator=g, device="cuda").bfloat16().transpose(1, 2)
            k = torch.randn(1, s, hkv, d, generator=g, device="cuda").bfloat16().transpose(1, 2)
        for _ in range(3):
            o = flash_attn_func(q, k, k, causal=True)
    else:
        t = sum(spec)
        q = torch.randn(t, hq, d, generator=g, device="cuda").bfloat16()
        k = torch.randn(t, hkv, d, generator=g, device="cuda").bfloat16()
        cu = [0]
        for n in spec:
            cu.append(cu[-1] + n)
        for _ in range(3):
            o = flash_attn_varlen_func(q, k, k, cu, cu, causal=True)
    torch.cuda.synchronize()
    _lib.check_device(0)
    print(name, "ok", float(o.float().abs().mean()))


if __name__ == "__main__":
    if len(sys.argv) > 1:
        run(sys.argv[1])
    else:
        for name in CASES:
            r = subprocess.run([sys.executable, __file__, name], capture_output=True, text=True)
            tail = (r.stdout + r.stderr).strip().splitlines()
            print(name, "rc", r.returncode, "|", " ".join(tail[-1:])[:200], flush=True)
