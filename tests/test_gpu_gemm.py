"""Prefill path (N > 8): the tcgen05 GEMM against the oracle.  All calls go through the C-ABI."""
import numpy as np
import pytest
from conftest import make_x, tol_c
from oracle_lib import F32, Q2_K, Q4_0, Q4_1, Q5_0, Q5_1, Q8_0, TYPE_NAMES, VEC_DOT, W_TYPES
from test_gpu_parity import dev, weights

pytestmark = pytest.mark.gpu
TC_TYPES = [Q4_0, Q4_1, Q5_0, Q5_1, Q8_0]



@pytest.fixture(params=["scaled", "exact"], autouse=True)
def gemm_mode(gpu, request):
    """Every test runs on both tensor-core kernels: the scaled-operand mainloop (default) and the exact
    per-block drain (lammcu_set_gemm_exact)."""
    lib = gpu.lib()
    old = lib.lammcu_get_gemm_exact()
    lib.lammcu_set_gemm_exact(1 if request.param == "exact" else 0)
    yield request.param
    lib.lammcu_set_gemm_exact(old)


GEMM_SHAPES = [  # (M, N, K)
    (128, 128, 256), (128, 9, 256), (1, 16, 512), (200, 130, 512), (257, 33, 1024), (33, 18, 4096),
    (384, 256, 2048), (130, 300, 768), (64, 512, 11008),
]


@pytest.mark.parametrize("t", TC_TYPES + [Q2_K, F32])
@pytest.mark.parametrize("m,n,k", GEMM_SHAPES)
def test_gemm_vs_oracle(gpu, oracle, t, m, n, k, gemm_mode):
    w = weights(oracle, t, m, k, seed=m + 5 * n + k)
    x = make_x(np.random.default_rng(n * 3 + k), n, k)
    want = oracle.mul_mat(t, w, x)
    gpu.lib().lammcu_launch_count_reset()
    wd = dev(w)
    got = gpu.mul_mat(t, wd if t != F32 else wd.view(__import__("torch").float32), dev(x)).cpu().numpy()
    launches = gpu.lib().lammcu_launch_count()
    if t in TC_TYPES:
        # scaled: fused quantize+prep, GEMM; exact: quantize, prep, GEMM
        assert launches == (2 if gemm_mode == "scaled" else 3)
    elif gemm_mode == "scaled" and k % 256 == 0:
        # Q2_K / F32 weights: (quantize +) weight dequantize + operand prep + TMA-fed tcgen05 GEMM
        assert launches == (4 if t == Q2_K else 3)
    err = np.abs(got - want).max()
    assert err <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{n}x{k}: err {err} tol {tol_c(want, k)}"


@pytest.mark.parametrize("t", TC_TYPES)
def test_gemm_matches_gemv_path(gpu, oracle, t, gemm_mode):
    """The tensor-core path and the dp4a GEMV path compute the same block sums: results agree to
    f32 summation-order noise (far inside the tolerance)."""
    m, n, k = 300, 40, 2048
    w = weights(oracle, t, m, k, seed=77)
    x = make_x(np.random.default_rng(78), n, k)
    a = gpu.mul_mat(t, dev(w), dev(x)).cpu().numpy()
    old = gpu.lib().lammcu_get_gemv_max_n()
    gpu.lib().lammcu_set_gemv_max_n(1 << 30)
    try:
        b = gpu.mul_mat(t, dev(w), dev(x)).cpu().numpy()
    finally:
        gpu.lib().lammcu_set_gemv_max_n(old)
    # exact kernel: only f32 summation order differs; scaled-operand kernel: + <= 2^-16 per product
    assert np.abs(a - b).max() <= (0.05 if gemm_mode == "exact" else 0.5) * tol_c(b, k)


@pytest.mark.parametrize("t", [Q4_1, Q8_0])
def test_gemm_llama_shapes_sampled(gpu, oracle, t):
    """BASELINE.json configs[2-3] shapes at N=512: oracle on a row sample, all 512 columns."""
    for (m, k) in [(4096, 4096), (11008, 4096)]:
        n = 512
        w = weights(oracle, t, m, k, seed=m + k + 1)
        rng = np.random.default_rng(3)
        x = rng.standard_normal((n, k)).astype(np.float32)
        got = gpu.mul_mat(t, dev(w), dev(x)).cpu().numpy()
        rows = np.unique(np.r_[0:4, 127:131, m - 4:m, rng.integers(0, m, 24)])
        want = oracle.mul_mat(t, w[rows], x)
        assert np.abs(got[:, rows] - want).max() <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{n}x{k}"
        # linearity in X (power-of-two scaling is exact through the quantizer)
        got2 = gpu.mul_mat(t, dev(w), dev(4 * x)).cpu().numpy()
        if t == Q8_0:
            assert np.array_equal(got2, 4 * got)


def test_gemm_strided_output_and_weights(gpu, oracle):
    import torch
    t, m, n, k = Q4_1, 150, 20, 512
    w = weights(oracle, t, m, k, seed=5)
    pad = np.zeros((m, w.shape[1] + 16), np.uint8)
    pad[:, : w.shape[1]] = w
    x = make_x(np.random.default_rng(6), n, k)
    want = oracle.mul_mat(t, w, x)
    out = torch.zeros((n, m + 7), dtype=torch.float32, device="cuda")
    gpu.mul_mat(t, dev(pad)[:, : w.shape[1]], dev(x), out=out[:, :m])
    got = out.cpu().numpy()
    assert np.abs(got[:, :m] - want).max() <= tol_c(want, k)
    assert (got[:, m:] == 0).all()
