"""the K-quants the reference lists as TODO (Q3_K, Q4_K, Q5_K, Q6_K x Q8_K) on the CUDA path,
against the oracle (pinned to ggml's own ggml_vec_dot_q*_K_q8_K, tests/test_oracle_vs_ref.py) and the golden
fixtures.  N == 1 runs the single-launch decode kernel (dot_unit_kx), 2 <= N <= 8 the streaming GEMV, prefill
the dequantize-once tcgen05 GEMM.  C-ABI only."""
import numpy as np
import pytest
from conftest import make_x, tol_c
from oracle_lib import KQ_TYPES, Q8_K, TYPE_NAMES, row_size
from test_gpu_parity import dev, weights

pytestmark = pytest.mark.gpu

GEMV_SHAPES = [  # (M, N, K)
    (33, 1, 4096), (1, 1, 256), (7, 3, 256), (130, 2, 512), (257, 5, 1024), (64, 8, 11008), (31, 7, 5120),
    (19, 4, 17920), (1000, 1, 2048),
]
GEMM_SHAPES = [(128, 128, 256), (128, 9, 256), (200, 130, 512), (257, 33, 1024), (33, 18, 4096), (64, 512, 11008)]


@pytest.mark.parametrize("t", KQ_TYPES)
@pytest.mark.parametrize("m,n,k", GEMV_SHAPES)
def test_kquant_gemv_vs_oracle(gpu, oracle, t, m, n, k):
    w = weights(oracle, t, m, k, seed=m * 7 + k + t)          # random bit patterns in every field
    x = make_x(np.random.default_rng(n + k), n, k)
    want, want_xq = oracle.mul_mat(t, w, x, want_xq=True)
    xq = gpu.quantize_act(dev(x), Q8_K).cpu().numpy()
    assert np.array_equal(xq, want_xq)
    got = gpu.mul_mat(t, dev(w), dev(x)).cpu().numpy()
    err = np.abs(got - want).max()
    assert err <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{n}x{k}: err {err} tol {tol_c(want, k)}"


@pytest.mark.parametrize("t", KQ_TYPES)
@pytest.mark.parametrize("m,n,k", GEMM_SHAPES)
def test_kquant_gemm_vs_oracle(gpu, oracle, t, m, n, k):
    w = weights(oracle, t, m, k, seed=m + 5 * n + k + t)
    x = make_x(np.random.default_rng(n * 3 + k), n, k)
    want = oracle.mul_mat(t, w, x)
    gpu.lib().lammcu_launch_count_reset()
    got = gpu.mul_mat(t, dev(w), dev(x)).cpu().numpy()
    # quantize + weight dequantize + operand prep + TMA-fed tcgen05 GEMM
    assert gpu.lib().lammcu_launch_count() == 4
    err = np.abs(got - want).max()
    assert err <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{n}x{k}: err {err} tol {tol_c(want, k)}"


@pytest.mark.parametrize("t", KQ_TYPES)
def test_kquant_llama7b_decode_shapes(gpu, oracle, t):
    """Full-size decode shapes on a row sample, plus exact power-of-two scaling of X (the Q8_K
    quantizer commutes with it bit for bit)."""
    for (m, k) in [(4096, 4096), (11008, 4096), (4096, 11008)]:
        w = weights(oracle, t, m, k, seed=m + k + t)
        rng = np.random.default_rng(23)
        x = rng.standard_normal((1, k)).astype(np.float32)
        got = gpu.mul_mat(t, dev(w), dev(x)).cpu().numpy()
        rows = np.r_[0:8, m // 2 - 3:m // 2 + 5, m - 8:m, rng.integers(0, m, 40)]
        want = oracle.mul_mat(t, w[rows], x)
        assert np.abs(got[:, rows] - want).max() <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{k}"
        assert np.array_equal(gpu.mul_mat(t, dev(w), dev(2 * x)).cpu().numpy(), 2 * got)


@pytest.mark.parametrize("t", KQ_TYPES)
def test_kquant_gemm_matches_gemv(gpu, oracle, t):
    """The same 24 columns through the prefill GEMM and 8 at a time through the GEMV."""
    m, n, k = 300, 24, 2048
    w, x = weights(oracle, t, m, k, seed=5 + t), make_x(np.random.default_rng(3), n, k)
    wd, xd = dev(w), dev(x)
    a = gpu.mul_mat(t, wd, xd).cpu().numpy()
    b = np.concatenate([gpu.mul_mat(t, wd, xd[i:i + 8].contiguous()).cpu().numpy() for i in range(0, n, 8)])
    want = oracle.mul_mat(t, w, x)
    assert np.abs(a - want).max() <= tol_c(want, k) and np.abs(b - want).max() <= tol_c(want, k)


@pytest.mark.parametrize("t", KQ_TYPES)
def test_kquant_operator(gpu, oracle, t):
    """lamm_can_mul_mat accepts the type (an extension of the reference's gate) and lamm_mul_mat on host
    ggml tensors gives the oracle's result; Xq written back to wdata is what ggml's INIT would have made."""
    m, n, k = 33, 18, 1024
    w = weights(oracle, t, m, k, seed=33 + t)
    x = make_x(np.random.default_rng(18), n, k)
    want, want_xq = oracle.mul_mat(t, w, x, want_xq=True)
    wdata = np.zeros((n, row_size(Q8_K, k)), np.uint8)
    gpu.lib().lamm_set_writeback_wdata(1)
    can, got = gpu.lamm_mul_mat_host(t, w, x, nth=4, wdata=wdata)
    gpu.lib().lamm_set_writeback_wdata(0)
    assert can
    assert np.abs(got - want).max() <= tol_c(want, k)
    assert np.array_equal(wdata, want_xq)
    can, one = gpu.lamm_mul_mat_host(t, w, x[:1], nth=1)       # N == 1 from host tensors
    assert can and np.abs(one - want[:1]).max() <= tol_c(want, k)
