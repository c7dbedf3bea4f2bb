"""scripts/benchmark_kernel.py: its in-script fp32 eager comparator must agree with the
oracle, and on a GPU the CLI must run end to end with the reference's flags."""
import importlib.util
import json
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def _load_cli():
    spec = importlib.util.spec_from_file_location("benchmark_kernel_cli", ROOT / "scripts" / "benchmark_kernel.py")
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


@pytest.mark.parametrize("causal", [False, True])
def test_cli_eager_matches_oracle(causal):
    from oracle.eager_attention import eager_attention

    cli = _load_cli()
    g = torch.Generator().manual_seed(3)
    q = torch.randn(2, 8, 33, 16, generator=g)
    k = torch.randn(2, 2, 40, 16, generator=g)
    v = torch.randn(2, 2, 40, 16, generator=g)
    a = cli.eager_attention(q, k, v, causal)
    b = eager_attention(q, k, v, None, causal)
    assert torch.allclose(a, b, atol=1e-6, rtol=0)


def test_cli_has_reference_flags():
    src = (ROOT / "scripts" / "benchmark_kernel.py").read_text()
    for flag, default in (("--batch-size", 16), ("--num-heads-q", 64), ("--num-heads-kv", 8), ("--seqlen-q", 1024),
                          ("--seqlen-kv", 1024), ("--dim", 128), ("--iter", 100)):
        assert f'"{flag}", type=int, default={default}' in src
    assert '"--dtype", type=str, default="half"' in src and '"--causal"' in src
    assert "import oracle" not in src and "from oracle" not in src


@pytest.mark.gpu
def test_cli_runs_on_gpu(tmp_path):
    out = tmp_path / "rec.json"
    r = subprocess.run([sys.executable, str(ROOT / "scripts" / "benchmark_kernel.py"), "--batch-size", "2",
                        "--num-heads-q", "8", "--num-heads-kv", "2", "--seqlen-q", "512", "--seqlen-kv", "512",
                        "--iter", "5", "--causal", "--dtype", "bfloat16", "--json", str(out)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    rec = json.loads(out.read_text())
    assert rec["accuracy"]["eager"]["max_abs"] <= 1e-2 and rec["accuracy"]["eager"]["rel_l2"] <= 5e-3
    assert rec["tflops"]["Custom"] > 0
