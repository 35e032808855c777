"""BASELINE.json's own configurations at FULL size (VERDICT r1 "parity holes"): every case runs the whole matrix on
the GPU through the C-ABI and is compared with the oracle on a sample of weight rows (first / last rows, tile and
shard boundaries, random rows) for all activation rows, plus size-independent properties (exact power-of-two
scaling through the quantizer; decode == the same columns of the N > 1 paths).

  config 3: Q2_K (super-block K-quant) decode GEMV and N = 512 prefill GEMM at the LLaMA-7B shapes
  config 4: LLaMA-13B Q4_1 prefill, N = 2048
  config 5: LLaMA-30B Q5_0 (K = 6656 / 17920: the K-split decode path, KS warps per row) at N = 1 and N = 512
  8f-1:     the K-quants beyond the reference's table at the largest decode shape (32000 x 4096)
"""
import numpy as np
import pytest
from conftest import tol_c
from oracle_lib import KQ_TYPES, Q2_K, Q4_1, Q5_0, TYPE_NAMES
from test_gpu_parity import dev

pytestmark = pytest.mark.gpu


class _ShapeOnly:
    """quantize_weights only needs .shape for the random-super-block formats: avoid materialising M x K floats."""
    def __init__(self, m, k):
        self.shape = (m, k)


def big_weights(oracle, t, m, k, seed):
    """32-element formats: N(0,1) through the reference quantizer, generated in row slabs so that the f32 source
    never exceeds ~0.25 GB; K-quants: random super-blocks (every scale / min / high-bit pattern)."""
    if t == Q2_K or t in KQ_TYPES:
        return oracle.quantize_weights(t, _ShapeOnly(m, k), seed=seed)
    rng = np.random.default_rng(seed)
    out = []
    step = max(1, (64 << 20) // k)
    for r0 in range(0, m, step):
        out.append(oracle.quantize_weights(t, rng.standard_normal((min(step, m - r0), k)).astype(np.float32)))
    return np.concatenate(out)


def sample_rows(m, rng, extra=()):
    rows = np.r_[0:6, 125:131, m // 2 - 2:m // 2 + 3, m - 6:m, rng.integers(0, m, 24), np.asarray(list(extra), np.int64)]
    return np.unique(np.clip(rows, 0, m - 1)).astype(np.int64)


def check_rows(oracle, t, w, x, got, k, rows, what):
    want = oracle.mul_mat(t, w[rows], x)
    err = np.abs(got[:, rows] - want).max()
    assert err <= tol_c(want, k), f"{what}: err {err} tol {tol_c(want, k)}"
    return want


@pytest.mark.parametrize("m,k", [(4096, 4096), (11008, 4096), (4096, 11008)])
def test_q2k_7b_decode_and_prefill_512(gpu, oracle, m, k):
    rng = np.random.default_rng(m + k)
    w = big_weights(oracle, Q2_K, m, k, seed=m ^ k)
    wd = dev(w)
    x = rng.standard_normal((512, k)).astype(np.float32)
    x[5] = 0.0
    x[6, : k // 2] = 0.0
    rows = sample_rows(m, rng)
    got = gpu.mul_mat(Q2_K, wd, dev(x)).cpu().numpy()
    check_rows(oracle, Q2_K, w, x, got, k, rows, f"q2_k {m}x512x{k}")
    got1 = gpu.mul_mat(Q2_K, wd, dev(x[:1])).cpu().numpy()                 # decode, all M rows vs the prefill's column 0
    check_rows(oracle, Q2_K, w, x[:1], got1, k, rows, f"q2_k {m}x1x{k}")
    assert np.abs(got1[0] - got[0]).max() <= 0.5 * tol_c(got[:1], k)
    assert np.array_equal(gpu.mul_mat(Q2_K, wd, dev(4 * x[:1])).cpu().numpy(), 4 * got1)


@pytest.mark.parametrize("m,k", [(13824, 5120), (5120, 13824), (5120, 5120)])
def test_q4_1_13b_prefill_2048(gpu, oracle, m, k):
    rng = np.random.default_rng(m + 3 * k)
    w = big_weights(oracle, Q4_1, m, k, seed=m + k)
    x = rng.standard_normal((2048, k)).astype(np.float32)
    x[1] = 0.0
    x[2, 0::7] = 2.5
    rows = sample_rows(m, rng, extra=[1727, 1728, 3455, 3456])            # 8-way shard boundaries of M = 13824
    got = gpu.mul_mat(Q4_1, dev(w), dev(x)).cpu().numpy()
    check_rows(oracle, Q4_1, w, x, got, k, rows, f"q4_1 {m}x2048x{k}")
    assert (got[1] == 0).all()


@pytest.mark.parametrize("m,k", [(17920, 6656), (6656, 17920), (6656, 6656)])
def test_q5_0_30b_decode_and_prefill_512(gpu, oracle, m, k):
    rng = np.random.default_rng(m + 5 * k)
    w = big_weights(oracle, Q5_0, m, k, seed=m + k)
    wd = dev(w)
    x = rng.standard_normal((512, k)).astype(np.float32)
    rows = sample_rows(m, rng, extra=[2239, 2240, 831, 832])               # 8-way shard boundaries
    got1 = gpu.mul_mat(Q5_0, wd, dev(x[:1])).cpu().numpy()                 # full M: the K-split decode path
    check_rows(oracle, Q5_0, w, x[:1], got1, k, rows, f"q5_0 {m}x1x{k}")
    assert np.array_equal(gpu.mul_mat(Q5_0, wd, dev(2 * x[:1])).cpu().numpy(), 2 * got1)
    got = gpu.mul_mat(Q5_0, wd, dev(x)).cpu().numpy()
    check_rows(oracle, Q5_0, w, x, got, k, rows, f"q5_0 {m}x512x{k}")
    assert np.abs(got1[0] - got[0]).max() <= 0.5 * tol_c(got[:1], k)       # decode vs the prefill's column 0
    got5 = gpu.mul_mat(Q5_0, wd, dev(x[:5])).cpu().numpy()                 # N = 2..8 kernel, same columns
    assert np.abs(got5 - got[:5]).max() <= 0.5 * tol_c(got[:5], k)


@pytest.mark.parametrize("t", KQ_TYPES + [Q2_K])
def test_kquants_largest_decode_shape(gpu, oracle, t):
    m, k = 32000, 4096
    rng = np.random.default_rng(t)
    w = big_weights(oracle, t, m, k, seed=t)
    x = rng.standard_normal((1, k)).astype(np.float32)
    got = gpu.mul_mat(t, dev(w), dev(x)).cpu().numpy()
    check_rows(oracle, t, w, x, got, k, sample_rows(m, rng), f"{TYPE_NAMES[t]} {m}x1x{k}")
