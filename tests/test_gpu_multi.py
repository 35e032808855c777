"""Multi-GPU parity inside the default `-m gpu` suite: spawns tests/multi_gpu_check.py under torchrun on every GPU the
box shows (2, 4 or 8; skipped on a 1-GPU box), so the driver's GPU test run sees the LL peer-store epilogue of the
decode kernel, the arrival-ticket tags, sharded sibling batches and the peer-store epilogue of the prefill GEMM.
The script checks every rank's gathered result against the oracle and exits non-zero on any mismatch."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_row_sharded_paths_under_torchrun(gpu):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU visible: the row-sharded paths need at least two (gpurun --gpus 2)")
    world = 8 if n >= 8 else (4 if n >= 4 else 2)
    gpu.lib().lammcu_host_quiesce()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29631", os.path.join(ROOT, "tests", "multi_gpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    tail = (r.stdout + r.stderr)[-4000:]
    assert r.returncode == 0, tail
    assert "FAIL" not in r.stdout, tail
    assert r.stdout.count(" ok") >= world, tail
