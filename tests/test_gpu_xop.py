"""SURVEY 8f-3 on the device: llm_build_norm (ggml_rms_norm + ggml_mul by the norm weight, llama.cpp:5171-5201) and
llm_build_ffn's silu(gate) * up (llama.cpp:5203-5290) fused into the activation quantizer / the decode kernel.
The oracle's restatement of both ops is bit-identical to stock ggml's graph (tests/test_oracle_vs_ref.py::
test_xops_vs_ggml_graph, the whole SiLU table included); here the CUDA path is compared with it: the quantized
activations byte for byte, the matmul within the tier tolerance, and the fixture stock ggml wrote."""
import numpy as np
import pytest
from conftest import golden, make_x, tol_c
from oracle_lib import F16, F32, KQ_TYPES, Q2_K, Q8_0, Q8_1, Q8_K, TYPE_NAMES, W_TYPES, vec_dot_of

pytestmark = pytest.mark.gpu
EPS = 1e-5


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def weights(oracle, t, m, k, seed):
    rng = np.random.default_rng(seed)
    if t == F32:
        return rng.standard_normal((m, k)).astype(np.float32).view(np.uint8).reshape(m, -1)
    if t == F16:
        return rng.standard_normal((m, k)).astype(np.float16).view(np.uint8).reshape(m, -1)
    return oracle.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=seed)


def inputs(kind, n, k, seed):
    """(x, aux) for the op: norm -> (x [n, k], norm weight [k]); silu -> (gate [n, k], up [n, k])"""
    rng = np.random.default_rng(seed)
    if kind == 1:
        return make_x(rng, n, k) * np.float32(3.0), (1.0 + 0.1 * rng.standard_normal(k)).astype(np.float32)
    gate = (rng.standard_normal((n, k)) * 4).astype(np.float32)
    gate[0, :6] = [0.0, -0.0, 1e-8, -1e-8, 20.0, -20.0]
    return gate, make_x(rng, n, k)


def op_of(gpu, kind, aux_d):
    return gpu.rms_norm_op(aux_d, EPS) if kind == 1 else gpu.silu_mul_op(aux_d)


def want_y(oracle, kind, x, aux):
    return oracle.rms_norm_mul(x, aux, EPS) if kind == 1 else oracle.silu_mul(x, aux)


def same_bits(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint8), np.ascontiguousarray(b).view(np.uint8))


@pytest.mark.parametrize("kind", [1, 2])
@pytest.mark.parametrize("vt", [Q8_0, Q8_1, Q8_K, F16, F32])
@pytest.mark.parametrize("n,k", [(1, 256), (1, 4096), (3, 11008), (17, 5120), (300, 4096)])
def test_op_quantize_bit_exact(gpu, oracle, kind, vt, n, k):
    x, aux = inputs(kind, n, k, seed=n * 7 + k + kind)
    y = want_y(oracle, kind, x, aux)
    got = gpu.quantize_act_xop(op_of(gpu, kind, dev(aux)), dev(x), vt).cpu().numpy()
    if vt == F32:
        assert same_bits(got, y)
    elif vt == F16:
        assert same_bits(got, y.astype(np.float16))
    else:
        assert np.array_equal(got, oracle.quantize_act(vt, y, prefill=0)), f"{(got != oracle.quantize_act(vt, y, prefill=0)).sum()} bytes differ"


def test_norm_without_weight_and_strided_rows(gpu, oracle):
    rng = np.random.default_rng(5)
    big = (rng.standard_normal((6, 1024 + 8)) * 2).astype(np.float32)
    xd = dev(big)[:, 8:]                                          # row stride 1032 floats, rows still 16-byte aligned
    y = oracle.rms_norm_mul(big[:, 8:], None, 1e-6)
    got = gpu.quantize_act_xop(gpu.rms_norm_op(None, 1e-6), xd, Q8_1).cpu().numpy()
    assert np.array_equal(got, oracle.quantize_act(Q8_1, y, prefill=0))
    gate, up = inputs(2, 6, 1024, 9)
    upb = np.zeros((6, 1024 + 12), np.float32)
    upb[:, :1024] = up
    got = gpu.quantize_act_xop(gpu.silu_mul_op(dev(upb)[:, :1024]), dev(gate), Q8_K).cpu().numpy()
    assert np.array_equal(got, oracle.quantize_act(Q8_K, oracle.silu_mul(gate, up), prefill=0))


def test_silu_table_every_fp16_pattern(gpu, oracle):
    """All 63 488 finite fp16 inputs: the device table is the one ggml_init builds (same host libm)."""
    h = np.arange(65536, dtype=np.uint16).view(np.float16).astype(np.float32)
    h = h[np.isfinite(h)].reshape(1, -1)
    h = h[:, : h.shape[1] // 8 * 8]
    one = np.ones_like(h)
    got = gpu.quantize_act_xop(gpu.silu_mul_op(dev(one)), dev(h), F32).cpu().numpy()
    assert same_bits(got, oracle.silu_mul(h, one))


@pytest.mark.parametrize("kind", [1, 2])
def test_golden_fixture(gpu, kind):
    """Fixture written by stock ggml's graph (oracle/gen_golden.py): y bit for bit, C within the tolerance."""
    g = golden("xop_q4_1_m33_n6_k2048.npz")
    k = g["x"].shape[1]
    x, aux, y, c = (g["x"], g["norm_w"], g["y_norm"], g["c_norm"]) if kind == 1 else (g["gate"], g["up"], g["y_silu"], g["c_silu"])
    op = op_of(gpu, kind, dev(aux))
    assert same_bits(gpu.quantize_act_xop(op, dev(x), F32).cpu().numpy(), y)
    got = gpu.mul_mat_xop(op, 3, dev(g["wq"]), dev(x)).cpu().numpy()
    assert np.abs(got - c).max() <= tol_c(c, k)
    got1 = gpu.mul_mat_xop(op_of(gpu, kind, dev(aux if kind == 1 else aux[2:3])), 3, dev(g["wq"]), dev(x[2:3])).cpu().numpy()
    assert np.abs(got1 - c[2:3]).max() <= tol_c(c, k)


@pytest.mark.parametrize("kind", [1, 2])
@pytest.mark.parametrize("t", W_TYPES + KQ_TYPES + [F16])
@pytest.mark.parametrize("m,n,k", [(515, 1, 4096), (1000, 1, 11008), (64, 5, 1024), (130, 40, 2048), (300, 200, 4096)])
def test_mul_mat_with_op(gpu, oracle, t, kind, m, n, k):
    """N == 1: one launch (the op inside the decode kernel), Xq written back and compared byte for byte;
    N <= 8: op+quantize then the streaming kernel; larger N: op+quantize then the prefill GEMM."""
    import torch
    w = weights(oracle, t, m, k, seed=m + k + t)
    x, aux = inputs(kind, n, k, seed=m + n + kind)
    y = want_y(oracle, kind, x, aux)
    want = oracle.mul_mat(t, w, y)
    wd, xd, op = dev(w), dev(x), op_of(gpu, kind, dev(aux))
    vt = vec_dot_of(t)
    ws = torch.zeros((n, 4 * k if t == F32 else gpu.row_size(vt, k)), dtype=torch.uint8, device="cuda")
    gpu.lib().lammcu_launch_count_reset()
    got = gpu.mul_mat_xop(op, t, wd, xd, ws=ws).cpu().numpy()
    launches = gpu.lib().lammcu_launch_count()
    err = np.abs(got - want).max()
    assert err <= tol_c(want, k), f"{TYPE_NAMES[t]} kind {kind} {m}x{n}x{k}: err {err} tol {tol_c(want, k)}"
    if n == 1 and t not in (F32, F16):
        assert launches == 1
    if t == F32:
        assert same_bits(ws.cpu().numpy(), y)
    elif t == F16:
        assert same_bits(ws.cpu().numpy(), y.astype(np.float16))
    else:
        assert np.array_equal(ws.cpu().numpy(), oracle.quantize_act(vt, y, prefill=0))
    # equals the unfused sequence on the device: op -> f32 -> mul_mat
    yd = gpu.quantize_act_xop(op, xd, F32).view(torch.float32)
    unfused = gpu.mul_mat(t, wd, yd).cpu().numpy()
    assert np.abs(got - unfused).max() <= 0.05 * tol_c(want, k)


@pytest.mark.parametrize("t", [3, 6, Q2_K, 12])
def test_sibling_batch_with_norm(gpu, oracle, t):
    """wq / wk / wv after attn_norm, gate / up after ffn_norm: ONE launch for norm + quantize + all matrices."""
    import torch
    k, ms = 4096, (4096, 1024, 1024)
    ws = [weights(oracle, t, m, k, seed=70 + j) for j, m in enumerate(ms)]
    x, nw = inputs(1, 1, k, seed=3)
    y = oracle.rms_norm_mul(x, nw, EPS)
    outs = [torch.zeros((1, m), dtype=torch.float32, device="cuda") for m in ms]
    op = gpu.rms_norm_op(dev(nw), EPS)
    gpu.lib().lammcu_launch_count_reset()
    gpu.mul_mat_batch_xop(op, t, [dev(w) for w in ws], dev(x), outs)
    assert gpu.lib().lammcu_launch_count() == 1
    for w, o in zip(ws, outs):
        want = oracle.mul_mat(t, w, y)
        assert np.abs(o.cpu().numpy() - want).max() <= tol_c(want, k)


def test_ffn_block_chain(gpu, oracle):
    """llm_build_ffn end to end at N = 1 in three launches: [ffn_norm + gate/up batch], [silu*mul + down]."""
    import torch
    t, k, f = 3, 2048, 5632
    wg, wu, wdn = weights(oracle, t, f, k, 1), weights(oracle, t, f, k, 2), weights(oracle, t, k, f, 3)
    x, nw = inputs(1, 1, k, seed=4)
    y = oracle.rms_norm_mul(x, nw, EPS)
    g, u = oracle.mul_mat(t, wg, y), oracle.mul_mat(t, wu, y)
    want = oracle.mul_mat(t, wdn, oracle.silu_mul(g, u))
    gd, ud = (torch.zeros((1, f), dtype=torch.float32, device="cuda") for _ in range(2))
    gpu.lib().lammcu_launch_count_reset()
    gpu.mul_mat_batch_xop(gpu.rms_norm_op(dev(nw), EPS), t, [dev(wg), dev(wu)], dev(x), [gd, ud])
    got = gpu.mul_mat_xop(gpu.silu_mul_op(ud), t, dev(wdn), gd).cpu().numpy()
    assert gpu.lib().lammcu_launch_count() == 2
    # gate / up carry the matmul tolerance into silu: compare against the oracle fed with the device's own g, u
    want_dev = oracle.mul_mat(t, wdn, oracle.silu_mul(gd.cpu().numpy(), ud.cpu().numpy()))
    assert np.abs(got - want_dev).max() <= tol_c(want_dev, f)
    assert np.abs(got - want).max() <= 50 * tol_c(want, f)


def test_bad_arguments_fail_loudly(gpu):
    import torch
    x = torch.zeros((1, 256), dtype=torch.float32, device="cuda")
    w = torch.zeros((4, 160), dtype=torch.uint8, device="cuda")
    with pytest.raises(gpu.LammError):
        gpu.mul_mat_xop(gpu.rms_norm_op(None, 0.0), 3, w, x)                 # eps must be > 0 (ggml asserts)
    bad = gpu.XOp(7, None, 0, 0.0)
    with pytest.raises(gpu.LammError):
        gpu.mul_mat_xop(bad, 3, w, x)
    up = torch.zeros((1, 260), dtype=torch.float32, device="cuda")[:, 1:257]    # rows not 16-byte aligned
    with pytest.raises(gpu.LammError):
        gpu.quantize_act_xop(gpu.silu_mul_op(up), x, Q8_0)
