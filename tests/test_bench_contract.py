"""bench.py's reference arm runs on host cores only, so its JSON contract is checked here on CPU: one line, the keys
the driver reads, the same metric / unit / config as the CUDA arm's, `impl`, a `cpu_baseline` describing the run and
an `e2e` object with zero transfer bytes.  (The CUDA arm needs a GPU; the driver runs it on a B200.)"""
import json
import os
import subprocess
import sys

from oracle_lib import ROOT, Ref


def test_reference_arm_prints_one_contract_line():
    if not Ref.available():
        import pytest
        pytest.skip("oracle/_ref/libref.so not built")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2",
                          "--warmup", "1"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference"
    assert d["metric"] == "llama7b_q4_1_decode_matmul_chain_tokens_per_s" and d["unit"] == "tok/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 1
    assert d["value"] > 0 and abs(d["value"] - 1e3 / d["ms_per_step"]) / d["value"] < 1e-6
    assert d["vs_baseline"] is None and d["data"] == "synthetic"
    assert d["config"]["workload"].startswith("LLaMA-7B Q4_1 decode") and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_cuda_arm_refuses_to_run_without_a_device():
    """No CPU fallback: without a GPU the default arm must fail loudly, not print a number."""
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a CUDA device is present")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "0", "--no-cpu"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode != 0
    assert not any(ln.lstrip().startswith("{") and '"value"' in ln for ln in out.stdout.splitlines())
