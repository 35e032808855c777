"""The oracle restatement against the golden fixtures produced by the real reference
(oracle/gen_golden.py).  CPU only."""
import glob
import os

import numpy as np
import pytest
from conftest import GOLDEN, golden, tol_c
from oracle_lib import F16, F32, Q8_0, Q8_1, Q8_K, TYPE_NAMES, VEC_DOT

MULMAT = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLDEN, "mulmat_*.npz")))


def test_fixtures_present():
    assert len(MULMAT) >= 8
    assert os.path.exists(os.path.join(GOLDEN, "quantize_act_n12_k2048.npz"))


@pytest.mark.parametrize("vt,key", [(Q8_0, "q8_0"), (Q8_1, "q8_1"), (Q8_K, "q8_k")])
def test_quantize_act_bit_exact(oracle, vt, key):
    g = golden("quantize_act_n12_k2048.npz")
    got = oracle.quantize_act(vt, g["x"], prefill=0xAB)
    assert np.array_equal(got, g[key]), f"{key}: {(got != g[key]).sum()} bytes differ"


def test_q8_K_zero_block_leaves_bsums(oracle):
    g = golden("quantize_act_n12_k2048.npz")
    q = g["q8_k"].reshape(12, 2048 // 256, 292)
    # row 1 is all zero: d == 0, qs == 0, bsums still hold the 0xAB pre-fill
    assert (q[1, :, :260] == 0).all() and (q[1, :, 260:] == 0xAB).all()


@pytest.mark.parametrize("name", MULMAT)
def test_mul_mat_vs_reference(oracle, name):
    g = golden(name)
    t = int(g["w_type"])
    x = g["x"]
    k = x.shape[1]
    if t == F32:
        c = oracle.mul_mat(t, g["wq"], x)
        assert np.abs(c - g["c_graph"]).max() <= tol_c(g["c_graph"], k)
        return
    # quantized activations: bit-exact (bytes the reference never writes are masked by pre-fill 0)
    c, xq = oracle.mul_mat(t, g["wq"], x, want_xq=True)
    ref_xq = g["xq"].copy()
    if VEC_DOT[t] == Q8_K:
        blk = ref_xq.reshape(x.shape[0], -1, 292)
        dead = (blk[:, :, :4] == 0).all(axis=2)
        blk[dead, 260:] = 0
    assert np.array_equal(xq, ref_xq)
    # C: ggml's own vec_dot is right for every format; lamm's kernels for all but Q8_0 on AVX2 (9-2)
    assert np.abs(c - g["c_vecdot"]).max() <= tol_c(g["c_vecdot"], k)
    if "c_lamm" not in g.files:                                # K-quants beyond Q2_K: lamm's gate refuses them,
        assert np.abs(c - g["c_graph"]).max() <= tol_c(g["c_graph"], k)   # ggml's own mul_mat took the graph
    elif TYPE_NAMES[t] != "q8_0":
        assert np.abs(c - g["c_lamm"]).max() <= tol_c(g["c_lamm"], k)
    # dequantizer agrees with ggml's to_float
    assert np.array_equal(oracle.dequantize(t, g["wq"], k), g["w_deq"])


@pytest.mark.parametrize("name", MULMAT)
def test_f32_order_variant(oracle, name):
    g = golden(name)
    t = int(g["w_type"])
    if t == F32:
        return
    k = g["x"].shape[1]
    xq = oracle.quantize_act(VEC_DOT[t], g["x"])
    a = oracle.mul_mat_q(t, g["wq"], xq, k)
    b = oracle.mul_mat_q(t, g["wq"], xq, k, f32_order=True)
    assert np.abs(a - b).max() <= 0.05 * tol_c(a, k)


def test_f16_weights_golden(oracle):
    """SURVEY 8f-2, oracle step: F16 weights against stock ggml's graph result (fixture from the compiled
    reference; the CUDA path does not take F16 yet)."""
    g = golden("next_f16_m33_n18_k4096.npz")
    assert int(g["w_type"]) == F16
    k = g["x"].shape[1]
    c = oracle.mul_mat(F16, g["wq"], g["x"])
    assert np.isfinite(c).all() and np.abs(c - g["c_graph"]).max() <= tol_c(g["c_graph"], k)
    assert np.array_equal(oracle.dequantize(F16, g["wq"], k), g["wq"].view(np.float16).astype(np.float32))
    assert np.array_equal(oracle.quantize_act(F16, g["x"]).view(np.uint16).reshape(g["x"].shape),
                          g["x"].astype(np.float16).view(np.uint16))


def test_xop_golden(oracle):
    """SURVEY 8f-3, oracle step: RMS norm * weight and silu(gate) * up against the fixture stock ggml's graph wrote:
    the f32 activations bit for bit, the Q4_1 matmul that consumes them within the tier tolerance."""
    g = golden("xop_q4_1_m33_n6_k2048.npz")
    k = g["x"].shape[1]
    y = oracle.rms_norm_mul(g["x"], g["norm_w"], float(g["eps"]))
    assert np.array_equal(y.view(np.uint32), g["y_norm"].view(np.uint32))
    assert np.abs(oracle.mul_mat(3, g["wq"], y) - g["c_norm"]).max() <= tol_c(g["c_norm"], k)
    y = oracle.silu_mul(g["gate"], g["up"])
    assert np.array_equal(y.view(np.uint32), g["y_silu"].view(np.uint32))
    assert np.abs(oracle.mul_mat(3, g["wq"], y) - g["c_silu"]).max() <= tol_c(g["c_silu"], k)


def test_threaded_port_matches(oracle):
    g = golden("mulmat_q4_1_m33_n18_k512.npz")
    a = oracle.mul_mat(3, g["wq"], g["x"])
    for nth in (1, 2, 4, 5):          # 5 does not divide 33: remainder rows still computed
        b = oracle.mul_mat_mt(3, g["wq"], g["x"], nth)
        assert np.abs(a - b).max() <= 0.05 * tol_c(a, 512)
