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
        # scaled: weight-row lift, fused quantize+prep, GEMM; exact: quantize, prep, GEMM
        assert launches == 3
    elif k % 256 == 0 and (t == Q2_K or gemm_mode == "scaled"):
        # Q2_K: quantize + int8 operand prep + kind::i8 GEMM; F32: weight split + operand prep + TMA-fed GEMM
        assert launches == 3
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


PAIR_FORMS = [("0", None), ("2", "0"), ("2", "1")]      # (LAMM_GEMM_PAIR, LAMM_GEMM_ATM)


@pytest.mark.parametrize("form", PAIR_FORMS, ids=["single", "pair-smem", "pair-tmem"])
@pytest.mark.parametrize("t", TC_TYPES)
def test_gemm_every_form_on_ragged_shapes(gpu, oracle, t, form, gemm_mode):
    """The scaled-operand GEMM has three forms -- single CTA, CTA pair with the unpacked weights in shared memory, CTA
    pair with them in tensor memory -- and the launcher picks by wave count, so small shapes only ever see the first.
    Both knobs are read per call: force each form on shapes whose last tiles are ragged in M and N (a pair whose peer
    CTA lies wholly beyond M; N tails of 1 and 255 columns; several tiles per pair so the rings and the accumulator
    hand-off wrap) and compare with the oracle; the forms must also agree with each other to summation-order noise."""
    import os
    if gemm_mode == "exact":
        pytest.skip("forms of the scaled-operand kernel")
    pair, atm = form
    old = {k_: os.environ.get(k_) for k_ in ("LAMM_GEMM_PAIR", "LAMM_GEMM_ATM")}
    os.environ["LAMM_GEMM_PAIR"] = pair
    if atm is not None:
        os.environ["LAMM_GEMM_ATM"] = atm
    try:
        # the last shape is 10 x 9 = 90 pair tiles on 74 pair slots: some pairs walk two tiles (oracle on a row sample)
        for (m, n, k) in [(300, 260, 512), (129, 129, 256), (257, 511, 1024), (640, 257, 768), (2500, 2100, 256)]:
            w = weights(oracle, t, m, k, seed=m + n)
            rng = np.random.default_rng(m * 7 + n)
            x = make_x(rng, n, k)
            rows = np.arange(m) if m < 1000 else np.unique(np.r_[0:3, 255:258, m - 3:m, rng.integers(0, m, 40)])
            want = oracle.mul_mat(t, w[rows], x)
            got = gpu.mul_mat(t, dev(w), dev(x)).cpu().numpy()[:, rows]
            err = np.abs(got - want).max()
            assert err <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{n}x{k} form {form}: err {err} tol {tol_c(want, k)}"
    finally:
        for k_, v in old.items():
            if v is None:
                os.environ.pop(k_, None)
            else:
                os.environ[k_] = v


def rescale_blocks(t, w, k, scale, rng=None):
    """Multiply the fp16 scale fields (d, and m / dmin where the format has one) of every block by `scale`, in
    fp16 arithmetic, leaving the quants alone; optionally plant one block per row with d = fp16 max and one with
    a subnormal d.  Returns new block bytes (the oracle and the GPU then see identical weights)."""
    from oracle_lib import TSIZE, BLCK
    bs, nblk = TSIZE[t], k // BLCK[t]
    blk = w.reshape(w.shape[0], nblk, bs).copy()
    offs = {Q4_0: [0], Q5_0: [0], Q8_0: [0], Q4_1: [0, 2], Q5_1: [0, 2], Q2_K: [80, 82]}[t]
    for o in offs:
        h = blk[:, :, o:o + 2].copy().view(np.float16)
        h = (h.astype(np.float32) * np.float32(scale)).astype(np.float16)
        blk[:, :, o:o + 2] = h.view(np.uint8)
    if rng is not None:
        rows = np.arange(w.shape[0])
        big = rng.integers(0, nblk, w.shape[0])
        tiny = (big + 1 + rng.integers(0, nblk - 1, w.shape[0])) % nblk
        blk[rows[::3], big[::3], offs[0]:offs[0] + 2] = np.array([65504.0], np.float16).view(np.uint8)
        blk[rows[1::3], tiny[1::3], offs[0]:offs[0] + 2] = np.array([3e-7], np.float16).view(np.uint8)   # subnormal
    return blk.reshape(w.shape[0], -1)


@pytest.mark.parametrize("t", TC_TYPES + [Q2_K])
@pytest.mark.parametrize("scale", [3e-5, 700.0, "planted"])
def test_gemm_weight_scale_sweep(gpu, oracle, t, scale, gemm_mode):
    """The reference's f32 path is range-safe for ANY fp16 block scale; so must the tensor-core paths be: weight
    scales down to the fp16 subnormals and up to fp16 max (a fixed pre-scale used to overflow the hi/lo split
    at |d q| 64 > 65504 and to lose the lo term for small d), rows that mix a huge and a tiny block."""
    m, n, k = 200, 40, 1024
    rng = np.random.default_rng(int(t) * 7 + 1)
    w = weights(oracle, t, m, k, seed=t + 11)
    w = rescale_blocks(t, w, k, 1.0, rng) if scale == "planted" else rescale_blocks(t, w, k, scale)
    x = make_x(rng, n, k)
    want = oracle.mul_mat(t, w, x)
    assert np.isfinite(want).all()
    got = gpu.mul_mat(t, dev(w), dev(x)).cpu().numpy()
    assert np.isfinite(got).all(), f"{TYPE_NAMES[t]} scale {scale}: non-finite C"
    # per weight row: a row's error is judged against that row's own magnitude (rows differ by 1e9 here)
    for i in range(m):
        tol = 1e-5 * np.sqrt(k) * max(float(np.abs(want[:, i]).max()), 1e-30)
        assert np.abs(got[:, i] - want[:, i]).max() <= tol, f"{TYPE_NAMES[t]} scale {scale} row {i}"


@pytest.mark.parametrize("scale", [1e-7, 1e4, 3e30, 1e-30])
def test_gemm_f32_weight_range(gpu, oracle, scale, gemm_mode):
    """F32 weights far outside fp16's range (|w| >= 1024 used to become inf in the hi/lo split)."""
    m, n, k = 150, 20, 512
    rng = np.random.default_rng(3)
    w = (rng.standard_normal((m, k)) * scale).astype(np.float32)
    w[::5] *= np.float32(1e-3)                                            # rows of very different magnitude
    x = make_x(rng, n, k)
    want = oracle.mul_mat(F32, w.view(np.uint8).reshape(m, -1), x)
    got = gpu.mul_mat(F32, dev(w), dev(x)).cpu().numpy()
    assert np.isfinite(got).all()
    for i in range(m):
        tol = 1e-5 * np.sqrt(k) * float(np.abs(want[:, i]).max())
        assert np.abs(got[:, i] - want[:, i]).max() <= tol, f"f32 scale {scale} row {i}"


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
