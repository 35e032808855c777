"""SURVEY 8f-2 on the device: F16 weights.  ggml rounds src1 to fp16 (ggml_fp32_to_fp16_row, ggml.c:354-372) and
runs ggml_vec_dot_f16 (ggml.c:1414-1457); the oracle's F16 leg is pinned to stock ggml's graph result by
tests/test_oracle_golden.py::test_f16_weights_golden and tests/test_oracle_vs_ref.py.  Here the CUDA path (the
converter, the N <= 8 streaming kernel and the single-MMA tcgen05 prefill kernel) is compared with it."""
import numpy as np
import pytest
from conftest import golden, make_x, tol_c
from oracle_lib import F16

pytestmark = pytest.mark.gpu


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def f16_weights(m, k, seed, scale=1.0):
    w = (np.random.default_rng(seed).standard_normal((m, k)) * scale).astype(np.float16)
    return w.view(np.uint8).reshape(m, 2 * k)


def test_type_table_has_f16(gpu):
    assert gpu.lib().lammcu_type_supported(F16) == 1
    assert gpu.blck_size(F16) == 1 and gpu.type_size(F16) == 2 and gpu.vec_dot_type(F16) == F16
    assert gpu.row_size(F16, 4096) == 8192


@pytest.mark.parametrize("n,k", [(1, 8), (1, 4096), (3, 11008), (18, 4096), (300, 5120)])
def test_convert_bit_exact(gpu, oracle, n, k):
    x = make_x(np.random.default_rng(n + k), n, k)
    for scale in (1.0, 3e-6, 700.0, 1e5):          # subnormal halves, and values beyond 65504 (-> inf, as ggml)
        xs = x * np.float32(scale)
        got = gpu.quantize_act(dev(xs), F16).cpu().numpy()
        with np.errstate(over="ignore"):
            want = xs.astype(np.float16).view(np.uint8).reshape(n, 2 * k)
        assert np.array_equal(got, want), f"n={n} k={k} scale={scale}: {(got != want).sum()} bytes differ"
        if scale != 1e5:
            assert np.array_equal(oracle.quantize_act(F16, xs), want)


def test_golden_fixture(gpu):
    """Fixture written by the compiled reference (stock ggml graph, F16 x F32 mul_mat at 33 x 18 x 4096)."""
    g = golden("next_f16_m33_n18_k4096.npz")
    k = g["x"].shape[1]
    c = gpu.mul_mat(F16, dev(g["wq"]), dev(g["x"])).cpu().numpy()
    assert np.abs(c - g["c_graph"]).max() <= tol_c(g["c_graph"], k)
    c1 = gpu.mul_mat(F16, dev(g["wq"]), dev(g["x"][:1])).cpu().numpy()
    assert np.abs(c1 - g["c_graph"][:1]).max() <= tol_c(g["c_graph"], k)


SHAPES = [(33, 1, 4096), (1, 1, 8), (7, 3, 264), (130, 2, 512), (257, 5, 1024), (64, 8, 11008), (19, 4, 17920),
          (33, 18, 4096), (100, 17, 2048), (300, 40, 520), (4096, 130, 4096), (1000, 300, 2048), (128, 512, 4096)]


@pytest.mark.parametrize("m,n,k", SHAPES)
def test_mul_mat_vs_oracle(gpu, oracle, m, n, k):
    w = f16_weights(m, k, seed=m + k)
    x = make_x(np.random.default_rng(n + k), n, k)
    want = oracle.mul_mat(F16, w, x)
    got = gpu.mul_mat(F16, dev(w), dev(x)).cpu().numpy()
    err = np.abs(got - want).max()
    assert err <= tol_c(want, k), f"f16 {m}x{n}x{k}: err {err} tol {tol_c(want, k)}"


def test_pair_kernel_walks_several_tiles(gpu, oracle):
    """The CTA-pair F16 kernel is persistent: 2500 x 2100 is 10 x 9 = 90 pair tiles on 74 pair slots, so some pairs walk
    two tiles (ring phases run on, the second tile lands in the other accumulator buffer, the drain of the first hands
    its buffer back); ragged in M (the last pair's peer CTA lies beyond M) and in N.  Oracle on a row sample; the whole
    result must also equal the single-CTA kernel's to summation-order noise."""
    import os
    m, n, k = 2500, 2100, 320
    w = f16_weights(m, k, seed=9)
    rng = np.random.default_rng(10)
    x = make_x(rng, n, k)
    rows = np.unique(np.r_[0:3, 127:130, 255:258, 2300:2310, m - 3:m, rng.integers(0, m, 40)])
    want = oracle.mul_mat(F16, w[rows], x)
    got = gpu.mul_mat(F16, dev(w), dev(x)).cpu().numpy()
    assert np.abs(got[:, rows] - want).max() <= tol_c(want, k)
    old = os.environ.get("LAMM_GEMM_PAIR")
    os.environ["LAMM_GEMM_PAIR"] = "0"
    try:
        single = gpu.mul_mat(F16, dev(w), dev(x)).cpu().numpy()
    finally:
        if old is None:
            os.environ.pop("LAMM_GEMM_PAIR", None)
        else:
            os.environ["LAMM_GEMM_PAIR"] = old
    assert np.abs(got - single).max() <= 0.05 * tol_c(want, k)


def test_prefill_kernel_is_the_tensor_core_one(gpu, oracle):
    """N >= the GEMM threshold with 16-byte aligned rows: convert + ONE tcgen05 kernel (2 launches); the products
    of two halves are exact in the f32 accumulator, so the result is far inside the tolerance."""
    m, n, k = 512, 256, 4096
    w = f16_weights(m, k, seed=5)
    x = make_x(np.random.default_rng(6), n, k)
    wd, xd = dev(w), dev(x)
    gpu.mul_mat(F16, wd, xd)
    gpu.lib().lammcu_launch_count_reset()
    got = gpu.mul_mat(F16, wd, xd).cpu().numpy()
    assert gpu.lib().lammcu_launch_count() == 2
    want = oracle.mul_mat(F16, w, x)
    assert np.abs(got - want).max() <= 0.05 * tol_c(want, k)


def test_strided_weights_and_scale_range(gpu, oracle):
    import torch
    m, n, k = 70, 33, 1024
    for scale in (1.0, 2.0 ** -12, 200.0):
        w = f16_weights(m, k, seed=9, scale=scale)
        pad = np.zeros((m, 2 * k + 16), np.uint8)
        pad[:, : 2 * k] = w
        x = make_x(np.random.default_rng(10), n, k)
        want = oracle.mul_mat(F16, w, x)
        out = torch.zeros((n, m + 3), dtype=torch.float32, device="cuda")
        gpu.mul_mat(F16, dev(pad)[:, : 2 * k], dev(x), out=out[:, :m])
        got = out.cpu().numpy()
        assert np.abs(got[:, :m] - want).max() <= tol_c(want, k), scale
        assert (got[:, m:] == 0).all()


def test_bad_shape_fails_loudly(gpu):
    import torch
    w = torch.zeros((4, 2 * 252), dtype=torch.uint8, device="cuda")
    with pytest.raises(gpu.LammError):
        gpu.mul_mat(F16, w, torch.zeros((1, 252), dtype=torch.float32, device="cuda"))    # K % 8


@pytest.mark.parametrize("n", [1, 18, 200])
def test_operator(gpu, oracle, n):
    """lamm_can_mul_mat + lamm_mul_mat on hand-built ggml tensors with an F16 src0 (the K / V-cache and lm-head
    type of many GGUF files), Xq written back to wdata as ggml's INIT phase would have left it."""
    m, k = 33, 4096
    w = f16_weights(m, k, seed=33)
    x = make_x(np.random.default_rng(18), max(n, 8), k)[:n]
    want, want_xq = oracle.mul_mat(F16, w, x, want_xq=True)
    wdata = np.zeros((n, 2 * k), np.uint8)
    gpu.lib().lamm_set_writeback_wdata(1)
    can, got = gpu.lamm_mul_mat_host(F16, w, x, nth=4, wdata=wdata)
    gpu.lib().lamm_set_writeback_wdata(0)
    assert can
    assert np.abs(got - want).max() <= tol_c(want, k)
    assert np.array_equal(wdata, want_xq)
    can, again = gpu.lamm_mul_mat_host(F16, w, x, nth=1)
    assert can and np.array_equal(again, got)
    can, _ = gpu.lamm_mul_mat_host(F16, w[:, : 2 * 4092], x[:, :4092].copy())
    assert not can                                                       # K % 8: ggml's own path takes it


def test_operator_mul_mat_id(gpu, oracle):
    rng = np.random.default_rng(301)
    m, k, n_as = 130, 1024, 4
    ws = [f16_weights(m, k, seed=60 + e) for e in range(n_as)]
    for n, ids in [(1, np.array([[2, 0]], np.int32)),
                   (11, np.stack([rng.integers(0, n_as - 1, 11), np.full(11, 1)], axis=1).astype(np.int32))]:
        x = make_x(rng, max(n, 8), k)[:n]
        for id_ in (0, 1):
            can, got = gpu.lamm_mul_mat_id_host(F16, ws, ids, id_, x)
            assert can
            want = oracle.mul_mat_id(F16, ws, ids, id_, x)
            assert np.abs(got - want).max() <= tol_c(want, k)
