"""Opt-in checks of tuning knobs that are compiled in but OFF by default and not yet measured on
hardware.  They run only with LAMM_TEST_EXPERIMENTAL=1 so that the default `-m gpu` run exercises
exactly what ships:

    LAMM_TEST_EXPERIMENTAL=1 python -m pytest tests/test_gpu_experimental.py -m gpu -q

LAMM_DECODE_RING=3|4 (lamm_decode.cu, launch_decode): a deeper per-warp TMA ring in the single-launch
decode kernel.  The knob is read once per process, so each setting runs in a child process; the child
must give bit-identical results to ring depth 2 (same per-row arithmetic, only the copy schedule
differs) and stay within the tier tolerance of the oracle."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.environ.get("LAMM_TEST_EXPERIMENTAL"), reason="opt-in: LAMM_TEST_EXPERIMENTAL=1")]

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import sys, os
sys.path.insert(0, os.path.join(%(root)r, "tests"))
sys.path.insert(0, %(root)r)
import numpy as np, torch
import lamm_b200 as L
from oracle_lib import Oracle, W_TYPES, KQ_TYPES
from test_gpu_parity import weights
L.init(0)
orc = Oracle()
out = {}
for t in list(W_TYPES) + list(KQ_TYPES):
    if t == 0:
        continue
    for (m, k) in [(33, 4096), (1185, 4096), (300, 11008), (2500, 1024), (7, 512), (4096, 4096)]:
        w = weights(orc, t, m, k, seed=m + 3 * k + t)
        x = np.random.default_rng(m * k).standard_normal((1, k)).astype(np.float32)
        c = L.mul_mat(t, torch.from_numpy(w).cuda(), torch.from_numpy(x).cuda()).cpu().numpy()
        out["%%d_%%d_%%d" %% (t, m, k)] = c
np.savez(sys.argv[1], **out)
"""


def _run(ring, path, pair=False):
    env = dict(os.environ)
    env.pop("LAMM_DECODE_RING", None)
    env.pop("LAMM_DECODE_PAIR", None)
    if ring != 2:
        env["LAMM_DECODE_RING"] = str(ring)
    if pair:
        env["LAMM_DECODE_PAIR"] = "1"
    subprocess.run([sys.executable, "-c", CHILD % {"root": ROOT}, path], check=True, env=env, timeout=600)
    return dict(np.load(path))


@pytest.mark.parametrize("ring", [3, 4])
def test_decode_ring_depth(tmp_path, oracle, ring):
    from conftest import tol_c
    from test_gpu_parity import weights
    base = _run(2, str(tmp_path / "r2.npz"))
    deep = _run(ring, str(tmp_path / f"r{ring}.npz"))
    assert base.keys() == deep.keys()
    for key in base:
        assert np.array_equal(base[key], deep[key]), f"ring {ring}: {key} differs from ring 2"
    # and ring 2 itself against the oracle on one case per format (the rest is tests/test_gpu_parity.py)
    for key in base:
        t, m, k = (int(v) for v in key.split("_"))
        if (m, k) != (33, 4096):
            continue
        w = weights(oracle, t, m, k, seed=m + 3 * k + t)
        x = np.random.default_rng(m * k).standard_normal((1, k)).astype(np.float32)
        want = oracle.mul_mat(t, w, x)
        assert np.abs(deep[key] - want).max() <= tol_c(want, k)


def test_decode_pair_rows(tmp_path):
    """LAMM_DECODE_PAIR=1 (Q4_0 / Q4_1: two rows share each read of the activation quants): same arithmetic per row,
    so every result must equal the default path bit for bit."""
    base = _run(2, str(tmp_path / "base.npz"))
    pair = _run(2, str(tmp_path / "pair.npz"), pair=True)
    for key in base:
        assert np.array_equal(base[key], pair[key]), f"pair: {key} differs"


@pytest.mark.parametrize("t", [3, 8, 10, 12, 0])
def test_operator_mul_mat_id(gpu, oracle, t):
    """lamm_can_mul_mat_id + lamm_mul_mat_id (written after the round's GPU budget was spent) against
    orc_mul_mat_id, which is pinned to stock ggml's ggml_mul_mat_id graph (tests/test_oracle_vs_ref.py): one
    token (decode), scattered tokens (gather / scatter), consecutive tokens per expert, an expert nobody picks."""
    from conftest import make_x, tol_c
    from test_gpu_parity import weights
    rng = np.random.default_rng(300 + t)
    m, k, n_as = 130, 1024, 4
    ws = [weights(oracle, t, m, k, seed=50 * t + e) for e in range(n_as)]
    for n, ids in [(1, np.array([[2, 0]], np.int32)),
                   (11, np.stack([rng.integers(0, n_as - 1, 11), np.full(11, 1)], axis=1).astype(np.int32)),
                   (9, np.stack([np.repeat([0, 1, 3], 3), np.repeat([3, 3, 0], 3)], axis=1).astype(np.int32))]:
        x = make_x(rng, max(n, 8), k)[:n]
        for id_ in (0, 1):
            can, got = gpu.lamm_mul_mat_id_host(t, ws, ids, id_, x)
            assert can
            want = oracle.mul_mat_id(t, ws, ids, id_, x)
            assert np.abs(got - want).max() <= tol_c(want, k), f"type {t} n {n} id {id_}"


def test_gguf_rows_to_device_and_mul_mat(gpu, oracle):
    """SURVEY 8f-4 on the device (the parser itself is CPU-tested in tests/test_gguf.py): rows of the committed
    GGUF fixture registered straight from the file mapping -- whole tensors and a two-way row shard -- then
    multiplied through lammcu_mul_mat_host and compared with the oracle on the same bytes."""
    import ctypes as C
    from conftest import tol_c
    from test_gguf import FIXTURE
    lib = gpu.lib()
    with gpu.Gguf(FIXTURE) as g:
        try:
            for name in ("blk.0.attn_q.weight", "blk.0.attn_k.weight", "blk.0.ffn_down.weight", "blk.0.ffn_up.weight",
                         "blk.0.ffn_gate.weight"):
                i = g.find(name)
                _, t, _, ne, raw = g.tensor(i)
                k, m = ne[0], ne[1]
                wq = raw.reshape(m, -1)
                x = np.random.default_rng(m + k).standard_normal((3, k)).astype(np.float32)
                want = oracle.mul_mat(t, wq, x)
                for lo, hi in [(0, m), (0, m // 2), (m // 2, m)]:
                    w = g.upload_rows(i, lo, hi)
                    c = np.zeros((3, m), np.float32)
                    rc = lib.lammcu_mul_mat_host(w, x.ctypes.data_as(C.POINTER(C.c_float)), 3, k,
                                                 c.ctypes.data_as(C.POINTER(C.c_float)), m, None)
                    assert rc == 0, lib.lammcu_last_error()
                    assert np.abs(c[:, lo:hi] - want[:, lo:hi]).max() <= tol_c(want, k), f"{name} rows {lo}:{hi}"
        finally:
            lib.lammcu_weight_forget_all()                    # before the mapping goes away
