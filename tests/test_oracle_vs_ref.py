"""Pin the oracle against the real reference compiled into oracle/_ref/libref.so.
Runs only where that library exists (this container; it also travels to the GPU box)."""
import numpy as np
import pytest
from conftest import make_x, tol_c
from oracle_lib import F16, F32, KQ_TYPES, Q2_K, Q8_0, Q8_1, Q8_K, TYPE_NAMES, VEC_DOT, W_TYPES


@pytest.mark.parametrize("vt", [Q8_0, Q8_1, Q8_K])
def test_act_quantizers_bit_exact(oracle, ref, vt):
    rng = np.random.default_rng(7)
    x = make_x(rng, 9, 4096)
    for scale in (1.0, 1e-4, 250.0):
        a = oracle.quantize_act(vt, x * np.float32(scale), prefill=0x5C)
        b = ref.quantize_act(vt, x * np.float32(scale), prefill=0x5C)
        assert np.array_equal(a, b)


@pytest.mark.parametrize("t", [t for t in W_TYPES if t not in (F32, Q2_K)])
def test_weight_quantizers_bit_exact(oracle, ref, t):
    rng = np.random.default_rng(11)
    w = rng.standard_normal((5, 1024)).astype(np.float32)
    assert np.array_equal(oracle.quantize_weights(t, w), ref.quantize_weights(t, w))


@pytest.mark.parametrize("t", [t for t in W_TYPES if t != F32])
def test_mul_mat_vs_lamm_and_vecdot(oracle, ref, t):
    rng = np.random.default_rng(1234)
    m, n, k = 37, 6, 4096
    w = rng.standard_normal((m, k)).astype(np.float32)
    x = make_x(rng, n, k)
    wq = ref.quantize_weights(t, w)
    xq = ref.quantize_act(VEC_DOT[t], x)
    c = oracle.mul_mat_q(t, wq, xq, k)
    tol = tol_c(c, k)
    assert np.abs(ref.vec_dot(t, wq, xq, k) - c).max() <= tol
    if TYPE_NAMES[t] != "q8_0":            # lamm's AVX2 Q8_0 kernel is wrong
        assert np.abs(ref.lamm_mul_mat(t, wq, x, xq) - c).max() <= tol
    else:
        assert np.abs(ref.lamm_mul_mat(t, wq, x, xq) - c).max() > 100 * tol


def test_q2k_random_blocks(oracle, ref):
    rng = np.random.default_rng(5)
    m, n, k = 8, 3, 2048
    wq = oracle.quantize_weights(Q2_K, np.zeros((m, k), np.float32), seed=99)
    x = rng.standard_normal((n, k)).astype(np.float32)
    xq = ref.quantize_act(Q8_K, x)
    c = oracle.mul_mat_q(Q2_K, wq, xq, k)
    assert np.isfinite(c).all()
    assert np.abs(ref.lamm_mul_mat(Q2_K, wq, x, xq) - c).max() <= tol_c(c, k)
    assert np.array_equal(oracle.dequantize(Q2_K, wq, k), ref.dequantize(Q2_K, wq, k))


@pytest.mark.parametrize("t", KQ_TYPES)
def test_kquants_vs_ggml_vec_dot(oracle, ref, t):
    """Q3_K..Q6_K are outside lamm's table (its gate refuses them), so the pin is ggml's own
    ggml_vec_dot_q*_K_q8_K and dequantize_row_q*_K -- on weights from ggml's quantizer AND on random
    bit patterns (every scale / high-bit combination)."""
    rng = np.random.default_rng(77 + t)
    m, n, k = 19, 5, 2048
    x = make_x(rng, n, k)
    xq = ref.quantize_act(Q8_K, x)
    for wq in (ref.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32)),
               oracle.quantize_weights(t, np.zeros((m, k), np.float32), seed=1000 + t)):
        c = oracle.mul_mat_q(t, wq, xq, k)
        assert np.isfinite(c).all()
        assert np.abs(ref.vec_dot(t, wq, xq, k) - c).max() <= tol_c(c, k)
        assert np.abs(ref.graph_mul_mat(t, wq, x) - c).max() <= tol_c(c, k)     # lamm refuses -> ggml's own path
        assert np.array_equal(oracle.dequantize(t, wq, k), ref.dequantize(t, wq, k))
    with pytest.raises(AssertionError):
        ref.lamm_mul_mat(t, wq, x, xq)                   # the reference operator itself refuses these types


def test_f16_weights_vs_ggml_graph(oracle, ref):
    """SURVEY 8f-2, oracle step: F16 weights.  lamm's gate refuses F16, so the behaviour to match is stock
    ggml: src1 -> ggml_fp32_to_fp16_row (RNE), ggml_vec_dot_f16.  The conversions must be byte-exact (also
    against numpy's IEEE float16), C within the tier tolerance; K need not be a multiple of anything."""
    rng = np.random.default_rng(4242)
    for (m, n, k) in [(33, 18, 4096), (7, 1, 100), (5, 3, 33), (19, 2, 11008)]:
        w = rng.standard_normal((m, k)).astype(np.float32)
        x = rng.standard_normal((n, k)).astype(np.float32)
        x[0, ::3] *= np.float32(1e-6)                  # fp16 subnormals
        x[-1, 1::5] *= np.float32(3e3)
        x[-1, 0] = np.float32(1e6)                     # beyond the fp16 range: +inf in the row, as in ggml
        wq = oracle.quantize_weights(F16, w)
        assert np.array_equal(wq, ref.quantize_weights(F16, w))
        assert np.array_equal(wq.view(np.float16).reshape(m, k), w.astype(np.float16))
        xq = oracle.quantize_act(F16, x)
        with np.errstate(over="ignore"):
            assert np.array_equal(xq.view(np.uint16).reshape(n, k), x.astype(np.float16).view(np.uint16))
        assert np.array_equal(oracle.dequantize(F16, wq, k), ref.dequantize(F16, wq, k))
        c = oracle.mul_mat(F16, wq, x)
        g = ref.graph_mul_mat(F16, wq, x)
        fin = np.isfinite(c)
        assert np.array_equal(fin, np.isfinite(g)) and fin[:-1].all() and not fin[-1].any()
        cf, gf = np.where(fin, c, 0), np.where(fin, g, 0)
        assert np.abs(cf - gf).max() <= tol_c(cf, k)
        v = ref.vec_dot(F16, wq, xq, k)
        assert np.abs(np.where(fin, v, 0) - cf).max() <= tol_c(cf, k)


@pytest.mark.parametrize("t", [3, 8, Q2_K, 12, F16, F32])
def test_mul_mat_id_vs_ggml_graph(oracle, ref, t):
    """SURVEY 8f-2, oracle step: ggml_mul_mat_id (expert selection per token) against stock ggml's graph,
    for a lamm format, Q8_0, both K-quant families, F16 and F32; every column of ids, experts that get
    no token, tokens that all pick one expert, 1 and 3 threads."""
    rng = np.random.default_rng(900 + t)
    m, n, k, n_as, n_ids = 37, 11, 1024, 4, 2
    x = make_x(rng, n, k)
    if t in (3, 8):
        # ggml's INIT phase runs the AVX2 quantize_row_q8_{0,1}, which rounds exact .5 ties to even where the
        # _reference quantizer rounds away: make_x's tie rows would differ by design, so these
        # formats get plain Gaussian rows here and the tie behaviour stays with test_act_quantizers_bit_exact
        x = rng.standard_normal((n, k)).astype(np.float32)
    wqs = []
    for e in range(n_as):
        w = rng.standard_normal((m, k)).astype(np.float32)
        if t == F32:
            wqs.append(w.copy().view(np.uint8).reshape(m, k * 4))
        elif t in (Q2_K, 12):
            wqs.append(ref.quantize_weights(t, w))
        else:
            wqs.append(oracle.quantize_weights(t, w))
    ids = rng.integers(0, n_as - 1, (n, n_ids)).astype(np.int32)      # expert n_as-1 never chosen in column 0
    ids[:, 1] = 2                                                     # column 1: every token picks expert 2
    ids[0, 0] = 0
    for id_ in range(n_ids):
        want = oracle.mul_mat_id(t, wqs, ids, id_, x)
        per_token = np.stack([oracle.mul_mat(t, wqs[ids[j, id_]], x[j:j + 1])[0] for j in range(n)])
        assert np.array_equal(want, per_token)
        for nth in (1, 3):
            got = ref.graph_mul_mat_id(t, wqs, ids, id_, x, nth=nth)
            assert np.abs(got - want).max() <= tol_c(want, k), f"{TYPE_NAMES[t]} id {id_} nth {nth}"
    bad = ids.copy()
    bad[3, 0] = n_as
    with pytest.raises(AssertionError):
        oracle.mul_mat_id(t, wqs, bad, 0, x)


def test_fp16_tables_all_65536_halves(oracle, ref):
    """fp16 -> f32 of every half bit pattern: reference table (through dequantize_row_q8_0 with
    qs = 1, ggml-impl.h:237-250) == oracle == IEEE (numpy).  NaNs compared by 'is NaN'."""
    bits = np.arange(65536, dtype=np.uint16)
    blocks = np.zeros((65536, 34), np.uint8)
    blocks[:, 0:2] = bits.view(np.uint8).reshape(-1, 2)
    blocks[:, 2] = 1                                           # qs[0] = 1 -> y[0] = d
    k = 32 * 32
    rows = blocks.reshape(65536 // 32, 32 * 34)
    got_ref = ref.dequantize(Q8_0, rows, k).reshape(65536, 32)[:, 0]
    got_orc = np.array([oracle.L.orc_fp16_to_fp32(int(b)) for b in bits], np.float32)
    ieee = bits.view(np.float16).astype(np.float32)
    nan = np.isnan(ieee)
    assert np.array_equal(np.isnan(got_ref), nan) and np.array_equal(np.isnan(got_orc), nan)
    assert np.array_equal(got_ref[~nan].view(np.uint32), ieee[~nan].view(np.uint32))
    assert np.array_equal(got_orc[~nan].view(np.uint32), ieee[~nan].view(np.uint32))


def test_fp32_to_fp16_rne(oracle):
    """f32 -> fp16 narrowing (ggml-impl.h:119-120 / :202-226) is IEEE round-to-nearest-even,
    subnormals and overflow included."""
    rng = np.random.default_rng(3)
    x = np.concatenate([rng.standard_normal(20000).astype(np.float32) * s for s in (1e-8, 1e-5, 1e-3, 1.0, 300.0, 7e4)])
    x = np.concatenate([x, np.array([0.0, -0.0, 65504.0, 65519.9, 65520.0, 5.9604645e-8, 2.9802322e-8, 6.1035156e-5], np.float32)])
    with np.errstate(over="ignore"):
        want = x.astype(np.float16).view(np.uint16)
    got = np.array([oracle.L.orc_fp32_to_fp16(float(v)) for v in x], np.uint16)
    assert np.array_equal(got, want)


def test_xops_vs_ggml_graph(oracle, ref):
    """SURVEY 8f-3, oracle step: llm_build_norm (ggml_rms_norm + ggml_mul) and llm_build_ffn's silu(gate) * up against
    stock ggml's graph.  The f32 activations must be BIT-identical (same element order in the double sum, the same
    libm expf behind the SiLU table), and the matmul that consumes them within the tier tolerance."""
    rng = np.random.default_rng(83)
    for (m, n, k, t) in [(33, 18, 4096, 3), (7, 1, 256, 6), (19, 5, 11008, Q2_K), (5, 3, 512, 2)]:     # not Q8_0: SURVEY 9-2
        wq = oracle.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=k)
        x = make_x(rng, n, k) * np.float32(3.0)
        nw = (1.0 + 0.1 * rng.standard_normal(k)).astype(np.float32)
        for eps in (1e-5, 1e-6):
            for w_ in (nw, None):
                c, y = ref.graph_xop_mul_mat(1, t, wq, x, w_, eps, nth=1)
                yo = oracle.rms_norm_mul(x, w_, eps)
                assert np.array_equal(y.view(np.uint32), yo.view(np.uint32))
                co = oracle.mul_mat(t, wq, yo)
                plain = [r for r in range(n) if r not in (3, 4)]     # make_x's tie rows: x86 ggml's INIT quantizer (AVX2)
                assert np.abs(c - co)[plain].max() <= tol_c(co, k)   # rounds halves to even, the path's one away
        gate = (rng.standard_normal((n, k)) * 4).astype(np.float32)
        gate[0, :8] = [0.0, -0.0, 1e-8, -1e-8, 70000.0, -70000.0, 20.0, -20.0]      # fp16 zero / subnormal / inf inputs
        up = rng.standard_normal((n, k)).astype(np.float32)
        c, y = ref.graph_xop_mul_mat(2, t, wq, gate, up, nth=1)     # nth > 1 drops remainder rows in the reference
        yo = oracle.silu_mul(gate, up)
        same = (y.view(np.uint32) == yo.view(np.uint32)) | (np.isnan(y) & np.isnan(yo))
        assert same.all()
        fin = np.isfinite(yo).all(axis=1)
        if fin.any():
            co = oracle.mul_mat(t, wq, yo[fin])
            assert np.abs(c[fin] - co).max() <= tol_c(co, k)


def test_silu_table_all_fp16_patterns(oracle, ref):
    """Every fp16 input pattern of the SiLU table through the real ggml_silu (one 65 536-element row)."""
    h = np.arange(65536, dtype=np.uint16).view(np.float16).astype(np.float32).reshape(1, -1)
    ok = np.isfinite(h[0])
    wq = oracle.quantize_weights(8, np.zeros((1, 65536), np.float32))
    _, y = ref.graph_xop_mul_mat(2, 8, wq, np.where(ok, h, 0).astype(np.float32), np.ones_like(h))
    yo = oracle.silu_mul(np.where(ok, h, 0).astype(np.float32), np.ones_like(h))
    assert np.array_equal(y.view(np.uint32), yo.view(np.uint32))
