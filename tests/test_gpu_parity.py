"""Parity of the CUDA path against the oracle and the golden fixtures.  All calls go through
the C-ABI (ctypes -> liblammcu.so); torch only owns the device buffers."""
import ctypes as C
import glob
import os

import numpy as np
import pytest
from conftest import GOLDEN, golden, make_x, tol_c
from oracle_lib import F32, KQ_TYPES, Q2_K, Q4_1, Q8_0, Q8_1, Q8_K, TYPE_NAMES, VEC_DOT, W_TYPES, row_size

pytestmark = pytest.mark.gpu
QTYPES = [t for t in W_TYPES if t != F32]
DTYPES = QTYPES + KQ_TYPES          # everything the single-launch decode kernel takes (SURVEY 8f-1 K-quants included)


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def zero_dead_bsums(xq, n):
    """Q8_K: bsums of all-zero super-blocks are unwritten in the reference, 0 on the GPU."""
    blk = xq.reshape(n, -1, 292).copy()
    dead = (blk[:, :, :4] == 0).all(axis=2)
    blk[dead, 260:] = 0
    return blk.reshape(n, -1)


def weights(oracle, t, m, k, seed):
    rng = np.random.default_rng(seed)
    if t == F32:
        return rng.standard_normal((m, k)).astype(np.float32)
    return oracle.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=seed)


# ---------------------------------------------------------------- activation quantize
@pytest.mark.parametrize("vt,key", [(Q8_0, "q8_0"), (Q8_1, "q8_1"), (Q8_K, "q8_k")])
def test_quantize_golden_bit_exact(gpu, vt, key):
    g = golden("quantize_act_n12_k2048.npz")
    got = gpu.quantize_act(dev(g["x"]), vt).cpu().numpy()
    want = g[key]
    if vt == Q8_K:
        want = zero_dead_bsums(np.where(want == 0xAB, want, want), 12)
        # pre-fill bytes only survive in dead bsums, which zero_dead_bsums cleared
    assert np.array_equal(got, want), f"{(got != want).sum()} bytes differ"


@pytest.mark.parametrize("vt", [Q8_0, Q8_1, Q8_K])
@pytest.mark.parametrize("n,k", [(1, 256), (1, 4096), (3, 11008), (8, 4096), (17, 5120), (512, 4096)])
def test_quantize_vs_oracle_bit_exact(gpu, oracle, vt, n, k):
    rng = np.random.default_rng(n * 131 + k)
    x = make_x(rng, n, k)
    for scale in (1.0, 3e-5, 700.0):
        xs = x * np.float32(scale)
        got = gpu.quantize_act(dev(xs), vt).cpu().numpy()
        want = oracle.quantize_act(vt, xs, prefill=0)
        assert np.array_equal(got, want), f"{TYPE_NAMES[vt]} n={n} k={k} scale={scale}: {(got != want).sum()} bytes differ"


def test_quantize_strided_and_unaligned_rows(gpu, oracle):
    import torch
    rng = np.random.default_rng(3)
    big = rng.standard_normal((5, 1024 + 3)).astype(np.float32)
    xd = dev(big)[:, 3:]                      # row stride 1027 floats, base 12 B off alignment
    for vt in (Q8_0, Q8_1, Q8_K):
        got = gpu.quantize_act(xd, vt).cpu().numpy()
        assert np.array_equal(got, oracle.quantize_act(vt, big[:, 3:], prefill=0))


# ---------------------------------------------------------------- mul_mat, golden
@pytest.mark.parametrize("name", sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLDEN, "mulmat_*.npz"))))
def test_mul_mat_golden(gpu, name):
    g = golden(name)
    t, x = int(g["w_type"]), g["x"]
    n, k = x.shape
    if t == F32:
        w = dev(g["wq"].view(np.float32).reshape(-1, k))
        c = gpu.mul_mat(t, w, dev(x)).cpu().numpy()
        assert np.abs(c - g["c_graph"]).max() <= tol_c(g["c_graph"], k)
        return
    xq = gpu.quantize_act(dev(x), VEC_DOT[t]).cpu().numpy()
    want_xq = zero_dead_bsums(g["xq"], n) if VEC_DOT[t] == Q8_K else g["xq"]
    assert np.array_equal(xq, want_xq)
    c = gpu.mul_mat(t, dev(g["wq"]), dev(x)).cpu().numpy()
    assert np.abs(c - g["c_vecdot"]).max() <= tol_c(g["c_vecdot"], k)
    if "c_lamm" not in g.files:          # Q3_K..Q6_K: the reference's gate refuses them, ggml's own path ran
        assert np.abs(c - g["c_graph"]).max() <= tol_c(g["c_graph"], k)
    elif TYPE_NAMES[t] != "q8_0":
        assert np.abs(c - g["c_lamm"]).max() <= tol_c(g["c_lamm"], k)


# ---------------------------------------------------------------- mul_mat vs oracle
SHAPES = [  # (M, N, K): odd M/N exercise tile edges like the reference's 33x18 case
    (33, 1, 4096), (33, 18, 4096), (1, 1, 256), (7, 3, 256), (130, 2, 512), (257, 5, 1024),
    (64, 8, 11008), (31, 7, 5120), (19, 4, 17920), (100, 17, 2048),
]


@pytest.mark.parametrize("t", W_TYPES)
@pytest.mark.parametrize("m,n,k", SHAPES)
def test_mul_mat_vs_oracle(gpu, oracle, t, m, n, k):
    w = weights(oracle, t, m, k, seed=m * 7 + k)
    x = make_x(np.random.default_rng(n + k), n, k)
    want = oracle.mul_mat(t, w if t != F32 else w.view(np.uint8).reshape(m, -1), x)
    got = gpu.mul_mat(t, dev(w), dev(x)).cpu().numpy()
    err = np.abs(got - want).max()
    assert err <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{n}x{k}: err {err} tol {tol_c(want, k)}"


@pytest.mark.parametrize("t", W_TYPES)
def test_llama7b_shapes_full_size(gpu, oracle, t):
    """BASELINE.json config[1]: K=4096, M in {4096, 11008}, N=1.  Oracle on a row sample
    (the full oracle would take minutes), plus linearity in X as a size-independent property."""
    import torch
    for (m, k) in [(4096, 4096), (11008, 4096), (4096, 11008)]:
        w = weights(oracle, t, m, k, seed=m + k)
        rng = np.random.default_rng(17)
        x = rng.standard_normal((1, k)).astype(np.float32)
        got = gpu.mul_mat(t, dev(w), dev(x)).cpu().numpy()
        rows = np.r_[0:8, m // 2 - 3:m // 2 + 5, m - 8:m, rng.integers(0, m, 40)]
        wv = w if t != F32 else w.view(np.uint8).reshape(m, -1)
        want = oracle.mul_mat(t, wv[rows], x)
        assert np.abs(got[:, rows] - want).max() <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{k}"
        # power-of-two scaling of X is exact through the quantizer: C(2x) == 2 C(x) bit for bit
        got2 = gpu.mul_mat(t, dev(w), dev(2 * x)).cpu().numpy()
        if t not in (3, 7):          # Q4_1/Q5_1 carry the fp16-rounded s: exact too unless s overflows
            assert np.array_equal(got2, 2 * got)
        else:
            assert np.abs(got2 - 2 * got).max() <= tol_c(want, k)


def test_strided_weights_and_output(gpu, oracle):
    """Row stride larger than the row (nb01 > row_size) and ldc > M."""
    import torch
    t, m, n, k = 3, 50, 3, 1024
    w = weights(oracle, t, m, k, seed=1)
    pad = np.zeros((m, w.shape[1] + 12), np.uint8)
    pad[:, : w.shape[1]] = w
    x = make_x(np.random.default_rng(2), n, k)
    want = oracle.mul_mat(t, w, x)
    wd = dev(pad)[:, : w.shape[1]]
    out = torch.zeros((n, m + 5), dtype=torch.float32, device="cuda")
    gpu.mul_mat(t, wd, dev(x), out=out[:, :m])
    got = out.cpu().numpy()
    assert np.abs(got[:, :m] - want).max() <= tol_c(want, k)
    assert (got[:, m:] == 0).all()


def test_zero_activations_give_zero(gpu, oracle):
    for t in QTYPES:
        w = weights(oracle, t, 40, 512, seed=9)
        c = gpu.mul_mat(t, dev(w), dev(np.zeros((2, 512), np.float32))).cpu().numpy()
        assert (c == 0).all(), TYPE_NAMES[t]


def test_bad_arguments_fail_loudly(gpu):
    import torch
    w = torch.zeros((4, 20 * 8), dtype=torch.uint8, device="cuda")
    x = torch.zeros((1, 250), dtype=torch.float32, device="cuda")
    with pytest.raises(gpu.LammError):
        gpu.mul_mat(3, w, x)                       # K % 32 != 0
    with pytest.raises(gpu.LammError):
        gpu.mul_mat(16, w, torch.zeros((1, 256), dtype=torch.float32, device="cuda"))   # IQ2_XXS: not on the path
    with pytest.raises(gpu.LammError):
        gpu.mul_mat(12, w, torch.zeros((1, 128), dtype=torch.float32, device="cuda"))   # Q4_K needs K % 256 == 0


# ---------------------------------------------------------------- the drop-in operator
@pytest.mark.parametrize("t", W_TYPES)
def test_lamm_mul_mat_operator(gpu, oracle, t):
    """lamm_can_mul_mat + lamm_mul_mat driven with hand-built ggml structs, as ggml.c:10858-10863
    does, at the reference's test shape (M=33, N=18, K=4096), 4 'threads'."""
    m, n, k = 33, 18, 4096
    w = weights(oracle, t, m, k, seed=33)
    x = make_x(np.random.default_rng(18), n, k)
    wv = w if t != F32 else w.view(np.uint8).reshape(m, -1)
    want, want_xq = oracle.mul_mat(t, wv, x, want_xq=True) if t != F32 else (oracle.mul_mat(t, wv, x), None)
    wdata = np.zeros((n, row_size(VEC_DOT[t], k)), np.uint8) if t != F32 else None
    gpu.lib().lamm_set_writeback_wdata(1)
    can, got = gpu.lamm_mul_mat_host(t, wv, x, nth=4, wdata=wdata)
    gpu.lib().lamm_set_writeback_wdata(0)
    assert can
    assert np.abs(got - want).max() <= tol_c(want, k)
    if t != F32:
        assert np.array_equal(wdata, want_xq), "GPU Xq != what ggml INIT would have written to wdata"
    # second call hits the weight registry (no re-upload) and gives the same bits
    can, again = gpu.lamm_mul_mat_host(t, wv, x, nth=1)
    assert can and np.array_equal(again, got)


@pytest.mark.parametrize("t", DTYPES)
def test_operator_decode_host_modes(gpu, oracle, t):
    """N == 1 through the operator in both modes of the host path (one launch + stream synchronize
    per call; jobs posted to the resident decode server): identical bits, fresh X each call (a
    stale staging row or tag would show), shapes of different row counts interleaved, and GPU work
    from elsewhere (torch) between calls, which has to wait for the server's idle exit."""
    import torch
    lib = gpu.lib()
    shapes = [(33, 4096), (1000, 1024), (4096, 4096), (515, 11008), (1, 256)]
    ws = [weights(oracle, t, m, k, seed=7 * m + k) for m, k in shapes]
    rng = np.random.default_rng(t)
    default_mode = lib.lammcu_get_host_mode()
    try:
        for rep in range(3):
            for (m, k), w in zip(shapes, ws):
                x = rng.standard_normal((1, k)).astype(np.float32)
                want = oracle.mul_mat(t, w, x)
                got = {}
                for mode in (0, 1):
                    assert lib.lammcu_set_host_mode(mode) == 0
                    can, got[mode] = gpu.lamm_mul_mat_host(t, w, x, nth=1)
                    assert can
                assert np.abs(got[0] - want).max() <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{k}"
                assert np.array_equal(got[1], got[0]), f"{TYPE_NAMES[t]} {m}x{k}"
        # a run of server jobs back to back (no mode switch in between), torch work interleaved
        assert lib.lammcu_set_host_mode(1) == 0
        z = torch.ones(1024, device="cuda")
        for rep in range(40):
            (m, k), w = shapes[rep % len(shapes)], ws[rep % len(shapes)]
            x = rng.standard_normal((1, k)).astype(np.float32)
            can, got1 = gpu.lamm_mul_mat_host(t, w, x, nth=1)
            want = oracle.mul_mat(t, w, x)
            assert can and np.abs(got1 - want).max() <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{k} job {rep}"
            if rep % 7 == 3:
                assert float((z * 2).sum().item()) == 2048.0
            if rep % 11 == 5:
                import time
                time.sleep(0.002)                    # longer than the idle period: the server exits and is relaunched
        assert lib.lammcu_host_quiesce() == 0
        # hostile timeouts: the server gives up after 1 us idle / at the first job boundary 10 us after its
        # launch, so nearly every call meets a server that has just left or is leaving (re-post + relaunch)
        assert lib.lammcu_set_server_timeouts(1, 10) == 0
        for rep in range(30):
            (m, k), w = shapes[rep % len(shapes)], ws[rep % len(shapes)]
            x = rng.standard_normal((1, k)).astype(np.float32)
            can, got1 = gpu.lamm_mul_mat_host(t, w, x, nth=1)
            want = oracle.mul_mat(t, w, x)
            assert can and np.abs(got1 - want).max() <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{k} job {rep} (short-lived server)"
    finally:
        lib.lammcu_set_server_timeouts(200, 5000)
        lib.lammcu_set_host_mode(default_mode)
    assert lib.lammcu_set_host_mode(2) != 0


@pytest.mark.parametrize("t", DTYPES)
def test_operator_decode_repeated_rows(gpu, oracle, t):
    """Host mode 1 sends an activation row once when consecutive calls carry bit-identical rows (ggml
    gives wq/wk/wv and gate/up the same src1): the later jobs reuse the quantized row the resident
    instance holds.  Results must equal mode 0 bit for bit for: the same row against several weights,
    a row changed IN PLACE (first element, last element), a row of another K in between, and all of it
    again with a server that dies after every job (the re-posted job must carry the row)."""
    lib = gpu.lib()
    shapes = [(4096, 4096), (33, 4096), (1000, 4096), (515, 11008)]
    ws = [weights(oracle, t, m, k, seed=11 * m + k) for m, k in shapes]
    rng = np.random.default_rng(100 + t)
    xa = rng.standard_normal((1, 4096)).astype(np.float32)
    xb = rng.standard_normal((1, 11008)).astype(np.float32)
    # (weight index, row, edit applied in place before the call); rows marked True may be reused
    plan = [(0, xa, None, False), (1, xa, None, True), (2, xa, None, True), (0, xa, None, True),
            (1, xa, 0, False), (2, xa, None, True), (0, xa, 4095, False), (3, xb, None, False),
            (3, xb, None, True), (1, xa, None, False), (2, xa, None, True)]
    default_mode = lib.lammcu_get_host_mode()
    prof = (C.c_double * 8)()
    try:
        def run(mode):
            xa0, xb0 = xa.copy(), xb.copy()
            assert lib.lammcu_set_host_mode(mode) == 0
            out = []
            for wi, x, edit, _ in plan:
                if edit is not None:
                    x[0, edit] += 1.5
                can, c = gpu.lamm_mul_mat_host(t, ws[wi], x, nth=1)
                assert can
                out.append(c)
            xa[:], xb[:] = xa0, xb0
            return out
        ref = run(0)
        want = oracle.mul_mat(t, ws[0], xa)
        assert np.abs(ref[0] - want).max() <= tol_c(want, 4096)
        assert not np.array_equal(ref[1], ref[4]) and not np.array_equal(ref[3], ref[6])  # the edits matter
        assert lib.lammcu_set_server_timeouts(100000, 4000000) == 0       # one instance for the whole run
        lib.lammcu_host_profile(prof, 1)
        got = run(1)
        lib.lammcu_host_profile(prof, 1)
        for i, (g, r) in enumerate(zip(got, ref)):
            assert np.array_equal(g, r), f"{TYPE_NAMES[t]} call {i}"
        expected = sum(1 for p_ in plan if p_[3])
        assert 1 <= prof[6] <= expected, (prof[6], expected)
        assert lib.lammcu_host_quiesce() == 0
        assert lib.lammcu_set_server_timeouts(1, 10) == 0                 # nearly every job meets a new instance
        for rep in range(3):
            got = run(1)
            for i, (g, r) in enumerate(zip(got, ref)):
                assert np.array_equal(g, r), f"{TYPE_NAMES[t]} call {i} (short-lived server, pass {rep})"
    finally:
        lib.lammcu_set_server_timeouts(200, 5000)
        lib.lammcu_set_host_mode(default_mode)


@pytest.mark.parametrize("t", [Q4_1, Q8_0, Q2_K])
def test_operator_sibling_speculation(gpu, oracle, t):
    """Host mode 1 learns which weight followed which with a bit-identical row (ggml: wq -> wk -> wv, gate -> up) and
    posts the followers' jobs behind the first one; a following call is answered from such a job only when it is for
    the predicted weight AND its row memcmp-equals the row the job used.  A decoder-layer-like call sequence over
    several tokens (fresh rows every token) must equal mode 0 bit for bit, with hits from the second token on; so must
    every way of breaking the prediction: a sibling skipped, a sibling's row edited in place, another weight in
    between, the same weight twice, weights forgotten and re-registered, and a server that dies after every job."""
    lib = gpu.lib()
    k, kf = 2048, 5632
    shapes = {"q": (512, k), "k": (136, k), "v": (136, k), "o": (512, k), "gate": (700, k), "up": (700, k), "down": (512, kf)}
    ws = {n: weights(oracle, t, m, kk, seed=17 * m + kk + i) for i, (n, (m, kk)) in enumerate(shapes.items())}
    rng = np.random.default_rng(7 + t)
    default_mode = lib.lammcu_get_host_mode()
    prof = (C.c_double * 8)()

    def token(rows, order, edit=None):
        out = []
        for j, name in enumerate(order):
            x = rows["f"] if name == "down" else rows[{"q": "a", "k": "a", "v": "a", "o": "b", "gate": "c", "up": "c"}[name]]
            if edit is not None and edit[0] == j:
                x = x.copy()
                x[0, edit[1]] += 0.75
            can, c = gpu.lamm_mul_mat_host(t, ws[name], x, nth=1)
            assert can
            out.append(c)
        return out
    layer = ["q", "k", "v", "o", "gate", "up", "down"]
    cases = [(layer, None)] * 4 + [
        (["q", "v", "k", "o", "gate", "down", "up"], None),            # siblings out of order, another weight in between
        (layer, None), (layer, (1, 0)), (layer, (2, k - 1)), (layer, (5, 7)),      # a sibling's row differs in one element
        (["q", "q", "k", "k", "v", "o", "gate", "gate", "up", "down"], None),   # the same weight twice in a row
        (layer, None), (layer, None)]
    toks = [{n: rng.standard_normal((1, kk)).astype(np.float32) for n, kk in (("a", k), ("b", k), ("c", k), ("f", kf))}
            for _ in cases]
    try:
        def run(mode):
            assert lib.lammcu_set_host_mode(mode) == 0
            return [token(rows, order, edit) for rows, (order, edit) in zip(toks, cases)]
        ref = run(0)
        want = oracle.mul_mat(t, ws["q"], toks[0]["a"])
        assert np.abs(ref[0][0] - want).max() <= tol_c(want, k)
        assert lib.lammcu_set_server_timeouts(100000, 4000000) == 0       # one instance for the whole run
        lib.lammcu_host_profile(prof, 1)
        got = run(1)
        lib.lammcu_host_profile(prof, 1)
        for i, (g, r) in enumerate(zip(got, ref)):
            for j, (a, b) in enumerate(zip(g, r)):
                assert np.array_equal(a, b), f"{TYPE_NAMES[t]} token {i} call {j}"
        assert prof[7] >= 3 * 3, f"speculative hits: {prof[7]} (k, v, up of tokens 2..4 at least)"
        # weights forgotten (their device copies freed) between tokens: every learned link must die with them
        assert lib.lammcu_weight_forget_all() == 0
        got = run(1)
        for i, (g, r) in enumerate(zip(got, ref)):
            for j, (a, b) in enumerate(zip(g, r)):
                assert np.array_equal(a, b), f"{TYPE_NAMES[t]} token {i} call {j} (after forget_all)"
        assert lib.lammcu_host_quiesce() == 0
        assert lib.lammcu_set_server_timeouts(1, 10) == 0                 # nearly every job meets a new instance
        for rep in range(2):
            got = run(1)
            for i, (g, r) in enumerate(zip(got, ref)):
                for j, (a, b) in enumerate(zip(g, r)):
                    assert np.array_equal(a, b), f"{TYPE_NAMES[t]} token {i} call {j} (short-lived server, pass {rep})"
    finally:
        lib.lammcu_set_server_timeouts(200, 5000)
        lib.lammcu_set_host_mode(default_mode)
        lib.lammcu_weight_forget_all()


def test_operator_tag_restart(gpu, oracle):
    """Job tags of the resident server are 32-bit; the host path restarts them from 1 (server stopped, every tagged
    buffer wiped) before they could wrap.  A child process whose tags start just below that point makes a few hundred
    host-tensor calls across it -- siblings, repeated rows and new rows -- and every result must equal the
    launch-per-call mode's bits."""
    import subprocess
    import sys
    child = r"""
import os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np
import lamm_b200 as L
from oracle_lib import Oracle, Q4_1
L.init(0)
lib = L.lib()
o = Oracle()
rng = np.random.default_rng(3)
ws = [o.quantize_weights(Q4_1, rng.standard_normal((m, 1024)).astype(np.float32), seed=m) for m in (300, 132, 132, 640)]
seq = [0, 1, 2, 3]
def run(mode, n):
    assert lib.lammcu_set_host_mode(mode) == 0
    r = np.random.default_rng(9)
    out = []
    for i in range(n):
        xa, xb = r.standard_normal((1, 1024)).astype(np.float32), r.standard_normal((1, 1024)).astype(np.float32)
        for wi in seq:
            can, c = L.lamm_mul_mat_host(Q4_1, ws[wi], xa if wi < 3 else xb, nth=1)
            assert can
            out.append(c)
    return out
ref = run(0, 60)
got = run(1, 60)
bad = [i for i, (a, b) in enumerate(zip(got, ref)) if not np.array_equal(a, b)]
print("calls", len(got), "bad", bad[:5])
sys.exit(1 if bad else 0)
""" % {"root": os.path.dirname(os.path.dirname(os.path.abspath(__file__)))}
    env = dict(os.environ, LAMM_SERVER_SEQ0=str(0xffffff00 - 100))
    r = subprocess.run([sys.executable, "-c", child], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


# ---------------------------------------------------------------- single-launch decode kernel (N == 1)
DECODE_SHAPES = [(33, 4096), (1, 256), (7, 512), (300, 11008), (64, 5120), (50, 17920), (9, 6656),
                 (1185, 4096), (2500, 1024), (129, 13824)]


@pytest.mark.parametrize("t", DTYPES)
@pytest.mark.parametrize("m,k", DECODE_SHAPES)
def test_decode_kernel_fused_quantize(gpu, oracle, t, m, k):
    """N == 1 through lammcu_mul_mat_dev: ONE launch (fused bit-exact activation quantize), Xq written
    to the caller's workspace must equal quantize_row_q8_*_reference, C within tolerance; the same
    operands through the two-launch generic path must agree."""
    import torch
    w = weights(oracle, t, m, k, seed=m + 3 * k)
    rng = np.random.default_rng(m * k)
    for variant in range(2):
        x = rng.standard_normal((1, k)).astype(np.float32)
        if variant == 1:                       # quantizer edge cases: zero blocks, ties, tiny values
            x[0, : k // 4] = 0.0
            x[0, k // 4: k // 4 + 64] = 2.5
            x[0, k // 2: k // 2 + 40] *= 1e-6
        want, want_xq = oracle.mul_mat(t, w, x, want_xq=True)
        ws = torch.full((row_size(VEC_DOT[t], k),), 0, dtype=torch.uint8, device="cuda")
        gpu.lib().lammcu_launch_count_reset()
        got = gpu.mul_mat(t, dev(w), dev(x), ws=ws).cpu().numpy()
        assert gpu.lib().lammcu_launch_count() == 1, "decode must be a single launch"
        assert np.abs(got - want).max() <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{k}"
        want_xq = zero_dead_bsums(want_xq, 1) if VEC_DOT[t] == Q8_K else want_xq
        assert np.array_equal(ws.cpu().numpy().reshape(1, -1), want_xq), "fused quantize is not bit-exact"
        # pre-quantized activations through the same kernel
        got_q = gpu.mul_mat_q(t, dev(w), dev(want_xq), k).cpu().numpy()
        assert np.array_equal(got_q, got), "fused and pre-quantized decode differ"
        # generic two-launch path
        gpu.lib().lammcu_set_decode_kernel(0)
        try:
            ref2 = gpu.mul_mat(t, dev(w), dev(x)).cpu().numpy()
        finally:
            gpu.lib().lammcu_set_decode_kernel(1)
        assert np.abs(ref2 - want).max() <= tol_c(want, k)


def test_decode_kernel_chain_pdl(gpu, oracle):
    """Back-to-back dependent decode launches (programmatic dependent launch): y = W2 (W1 x).
    The second kernel prefetches W2 early but must read W1's output only after it is complete."""
    import torch
    t, k = 3, 4096
    w1 = weights(oracle, t, 4096, k, seed=11)
    w2 = weights(oracle, t, 512, k, seed=12)
    x = np.random.default_rng(5).standard_normal((1, k)).astype(np.float32) * 0.05
    y1 = oracle.mul_mat(t, w1, x)
    y2 = oracle.mul_mat(t, w2, y1)
    w1d, w2d, xd = dev(w1), dev(w2), dev(x)
    for _ in range(20):
        a = gpu.mul_mat(t, w1d, xd)
        b = gpu.mul_mat(t, w2d, a)
        assert np.abs(b.cpu().numpy() - y2).max() <= 2 * tol_c(y2, k)


@pytest.mark.parametrize("t", DTYPES)
@pytest.mark.parametrize("k,ms", [(4096, (4096, 1024, 516)), (11008, (256, 1000)), (512, (33, 7, 64, 129))])
def test_decode_batch_siblings(gpu, oracle, t, k, ms):
    """lammcu_mul_mat_batch_dev: sibling matmuls (q/k/v, gate/up) that share one activation row in ONE
    launch.  Must equal the separate launches bit for bit (same per-row arithmetic) and the oracle within
    tolerance; shapes the fused launch cannot take (Q2_K rows whose prefix breaks 16-byte congruence)
    fall back to separate launches and still agree."""
    import torch
    rng = np.random.default_rng(k + sum(ms))
    x = make_x(rng, 1, k)
    ws = [weights(oracle, t, m, k, seed=100 + j) for j, m in enumerate(ms)]
    wd = [dev(w) for w in ws]
    xd = dev(x)
    outs = [torch.full((1, m), float("nan"), device="cuda") for m in ms]
    lib = gpu.lib()
    lib.lammcu_launch_count_reset()
    gpu.mul_mat_batch(t, wd, xd, outs)
    launches = int(lib.lammcu_launch_count())
    torch.cuda.synchronize()
    stride = ws[0].shape[1]
    prefix_ok = all((sum(ms[:j]) * stride) % 16 == 0 for j in range(len(ms)))
    if k % 256 == 0 and prefix_ok:
        assert launches == 1, f"expected one fused launch, saw {launches}"
    for w, wdev, o, m in zip(ws, wd, outs, ms):
        got = o.cpu().numpy()
        sep = gpu.mul_mat(t, wdev, xd).cpu().numpy()
        want = oracle.mul_mat(t, w, x)
        assert np.array_equal(got, sep), f"batch differs from separate launch (M={m})"
        assert np.abs(got - want).max() <= tol_c(want, k)


@pytest.mark.parametrize("t", [3, 8, 10, 12, 0])
def test_operator_mul_mat_id(gpu, oracle, t):
    """lamm_can_mul_mat_id + lamm_mul_mat_id against
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
