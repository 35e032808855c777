"""lammcu_chain_* (include/lammcu.h): a list of dependent N == 1 mul_mats -- a decoded token's q/k/v, wo, gate/up,
down, ..., head -- executed by ONE persistent launch.  The per-row arithmetic is the decode kernel's, so every
intermediate and final vector must equal, bit for bit, what one lammcu_mul_mat_dev / lammcu_mul_mat_batch_dev call
per op gives; the last vector is also checked against the oracle evaluated along the same chain."""
import numpy as np
import pytest
from conftest import tol_c
from oracle_lib import KQ_TYPES, TYPE_NAMES, W_TYPES
from test_gpu_parity import dev, weights

pytestmark = pytest.mark.gpu
DTYPES = [t for t in W_TYPES if t != 0] + KQ_TYPES


def build(gpu, oracle, t, H, F, V, layers, seed):
    import torch
    rng = np.random.default_rng(seed)

    def w(m, k, s):
        # unit-gain-ish weights keep the chained values finite: scale the activations instead (below)
        return weights(oracle, t, m, k, seed=s)
    x0 = dev(rng.standard_normal((1, H)).astype(np.float32))
    mk = lambda m: torch.zeros((1, m), device="cuda")
    ops, host = [], []
    xin = x0
    for layer in range(layers):
        q, k, v, o, g, u, xo = mk(H), mk(H), mk(H), mk(H), mk(F), mk(F), mk(H)
        wq, wk, wv, wo = (w(H, H, 10 * layer + i) for i in range(4))
        wg, wu, wd = w(F, H, 10 * layer + 4), w(F, H, 10 * layer + 5), w(H, F, 10 * layer + 6)
        for ws, x, outs in (([wq, wk, wv], xin, [q, k, v]), ([wo], v, [o]), ([wg, wu], o, [g, u]), ([wd], u, [xo])):
            host.append((ws, x, outs))
            ops.append(([dev(a) for a in ws], x, outs))
        xin = xo
    wh = w(V, H, 999)
    logits = mk(V)
    host.append(([wh], xin, [logits]))
    ops.append(([dev(wh)], xin, [logits]))
    return x0, ops, host, logits


@pytest.mark.parametrize("t", DTYPES)
def test_chain_equals_one_call_per_op(gpu, oracle, t):
    import torch
    # H = 1024 (one K segment), F = 5120 (the K-split path: two warps per row for the 32-element formats), ragged V
    H, F, V = 1024, 5120, 2001 if t not in (10,) + tuple(KQ_TYPES) else 2000
    x0, ops, host, logits = build(gpu, oracle, t, H, F, V, layers=2, seed=t)
    # values grow by ~sqrt(K) per matmul with N(0,1) weights: renormalise x0 so that nothing overflows in 9 stages
    x0.mul_(1e-12)
    ch = gpu.Chain(t, ops)
    try:
        for rep in range(2):                                  # twice: the barrier counter is reset per launch
            for _, _, outs in ops:
                for o in outs:
                    o.fill_(float("nan"))
            ch.run()
            torch.cuda.synchronize()
            assert not ch.timed_out()
            got = [[o.clone() for o in outs] for _, _, outs in ops]
            # the same list, one call per op, on the same buffers
            for ws, x, outs in ops:
                if len(ws) == 1:
                    gpu.mul_mat(t, ws[0], x, out=outs[0])
                else:
                    gpu.mul_mat_batch(t, ws, x, outs)
            torch.cuda.synchronize()
            for i, (_, _, outs) in enumerate(ops):
                for j, o in enumerate(outs):
                    assert torch.isfinite(o).all()
                    assert torch.equal(got[i][j], o), f"{TYPE_NAMES[t]} rep {rep} op {i} output {j} differs"
        # and against the oracle along the chain (each op on the GPU's own previous outputs: per-op tolerance)
        for (ws, x, outs) in host[-3:]:
            xh = x.cpu().numpy()
            for wj, oj in zip(ws, outs):
                want = oracle.mul_mat(t, wj, xh)
                assert np.abs(oj.cpu().numpy() - want).max() <= tol_c(want, xh.shape[1])
    finally:
        ch.close()


def test_chain_under_cuda_graph_and_bad_arguments(gpu, oracle):
    import ctypes as C
    import torch
    t = 3
    x0, ops, host, logits = build(gpu, oracle, t, 1024, 5120, 1000, layers=1, seed=5)
    x0.mul_(1e-6)
    ch = gpu.Chain(t, ops)
    try:
        ch.run()
        torch.cuda.synchronize()
        want = logits.clone()
        logits.zero_()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(s):
            ch.run()
        torch.cuda.synchronize()
        with torch.cuda.graph(g):
            ch.run()
        logits.zero_()
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        assert torch.equal(logits, want) and not ch.timed_out()
    finally:
        ch.close()
    lib = gpu.lib()
    h = C.c_void_p()
    bad = (gpu.ChainOp * 1)()
    bad[0].count = 5
    assert lib.lammcu_chain_create(t, 1, bad, C.byref(h)) != 0
    assert lib.lammcu_chain_create(0, 1, bad, C.byref(h)) == -3          # F32 weights: LAMMCU_ERR_TYPE
