"""The host operator's own logic on CPU (csrc/lamm_host.cpp: gates, batch planes and broadcast factors, strides,
pre-quantized src1, the wdata write-back hook, mul_mat_id grouping / gather / scatter), with the device half of the
C-ABI replaced by tests/hoststub/host_stub.cpp -- a stand-in built on the CPU oracle.  Test infrastructure only: the
stub library is built here, into tests/hoststub/, and is never part of the product.  What runs underneath on a GPU box
(lammcu_mul_mat_host etc.) has its own parity tests in tests/test_gpu_*.py."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
from oracle_lib import F16, F32, Q2_K, Q8_0, Q8_1, Q8_K, ROOT, TYPE_NAMES, VEC_DOT, row_size

HERE = os.path.join(ROOT, "tests", "hoststub")
LIB = os.path.join(HERE, "libhoststub.so")
I32 = 26


@pytest.fixture(scope="module")
def host(lamm):
    srcs = [os.path.join(HERE, "host_stub.cpp"), os.path.join(ROOT, "la-llama.cpp_b200", "csrc", "lamm_host.cpp"),
            os.path.join(ROOT, "oracle", "lamm_oracle.c"), os.path.join(ROOT, "include", "lammcu.h")]
    if not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs):
        obj = os.path.join(HERE, "lamm_oracle.o")
        subprocess.check_call(["gcc", "-std=c11", "-O2", "-ffp-contract=off", "-fPIC", "-c", srcs[2], "-o", obj])
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", srcs[0], srcs[1], obj, "-o", LIB,
                               "-lpthread", "-lm"])
    L = C.CDLL(LIB)
    P, T = C.POINTER(lamm.GgmlComputeParams), C.POINTER(lamm.GgmlTensor)
    for name in ("lamm_can_mul_mat", "lamm_can_mul_mat_id"):
        getattr(L, name).restype = C.c_bool
        getattr(L, name).argtypes = [P, T]
    for name in ("lamm_mul_mat", "lamm_mul_mat_id"):
        getattr(L, name).restype = None
        getattr(L, name).argtypes = [P, T]
    L.stub_calls.restype = C.c_long
    L.stub_rows.restype = C.c_long
    L.stub_refreshes.restype = C.c_long
    L.stub_reset()
    return L


def tensor(lamm, ttype, ne, nb, arr):
    t = lamm.GgmlTensor()
    t.type = ttype
    for i in range(4):
        t.ne[i] = ne[i] if i < len(ne) else 1
        t.nb[i] = nb[i] if i < len(nb) else nb[-1] * (ne[len(nb) - 1] if i == len(nb) else 1)
    for i in range(len(nb), 4):
        t.nb[i] = t.nb[i - 1] * t.ne[i - 1]
    t.data = arr.ctypes.data
    return t


def run(lamm, host, dst, nth=1, task=None, wdata=None, id_op=False):
    can_fn, mm_fn = (host.lamm_can_mul_mat_id, host.lamm_mul_mat_id) if id_op else (host.lamm_can_mul_mat, host.lamm_mul_mat)
    for ith in range(nth):
        p = lamm.GgmlComputeParams(lamm.TASK_COMPUTE if task is None else task, ith, nth,
                                   wdata.nbytes if wdata is not None else 0, wdata.ctypes.data if wdata is not None else None)
        if not can_fn(C.byref(p), C.byref(dst)):
            return False
        mm_fn(C.byref(p), C.byref(dst))
    return True


@pytest.mark.parametrize("t", [3, 6, Q8_0, Q2_K, 12, F32])
def test_mul_mat_planes_broadcast_and_strides(lamm, host, oracle, t):
    """ne02 = 2 weight planes broadcast over ne12 = 4 activation planes (r2 = 2), ne13 = 2 over ne03 = 1 (r3 = 2),
    padded row strides on src1 and dst, 3 threads (only ith == 0 works): every plane equals the oracle bit for bit."""
    rng = np.random.default_rng(t)
    m, n, k = 9, 3, 512
    ne02, ne12, ne13 = 2, 4, 2
    rs = row_size(t, k)
    w = np.stack([oracle.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=e).reshape(m, -1).view(np.uint8)
                  if t != F32 else rng.standard_normal((m, k)).astype(np.float32).view(np.uint8).reshape(m, -1)
                  for e in range(ne02)])
    # F32 weights pair with F32 activations, and a src1 that already has the vec-dot type must be contiguous
    # (src/loongarch_matmul.cpp:26): no row padding there
    ldx, ldc = (k if t == F32 else k + 16), m + 5
    x = rng.standard_normal((ne13, ne12, n, ldx)).astype(np.float32)
    c = np.full((ne13, ne12, n, ldc), 7.0, np.float32)
    src0 = tensor(lamm, t, [k, m, ne02, 1], [lamm.type_size(t), rs, rs * m], w)
    src1 = tensor(lamm, F32, [k, n, ne12, ne13], [4, ldx * 4, n * ldx * 4, ne12 * n * ldx * 4], x)
    dst = tensor(lamm, F32, [m, n, ne12, ne13], [4, ldc * 4, n * ldc * 4, ne12 * n * ldc * 4], c)
    dst.src[0], dst.src[1] = C.pointer(src0), C.pointer(src1)
    host.stub_reset()
    assert run(lamm, host, dst, nth=3)
    assert host.stub_calls() == ne12 * ne13 and host.stub_rows() == ne12 * ne13 * n
    for i13 in range(ne13):
        for i12 in range(ne12):
            want = oracle.mul_mat(t, w[i12 // 2], x[i13, i12, :, :k])
            assert np.array_equal(c[i13, i12, :, :m], want), f"{TYPE_NAMES.get(t, t)} plane {i13},{i12}"
    assert (c[..., m:] == 7.0).all()                                   # the padding of dst rows is untouched
    if t == F32:                                                       # the same node with padded src1 rows is refused
        xp = rng.standard_normal((ne13, ne12, n, k + 16)).astype(np.float32)
        src1p = tensor(lamm, F32, [k, n, ne12, ne13], [4, (k + 16) * 4, n * (k + 16) * 4, ne12 * n * (k + 16) * 4], xp)
        dst.src[1] = C.pointer(src1p)
        assert not run(lamm, host, dst)


def test_gates(lamm, host, oracle):
    rng = np.random.default_rng(0)
    m, n, k = 4, 2, 256
    w = oracle.quantize_weights(3, rng.standard_normal((m, k)).astype(np.float32))
    x = rng.standard_normal((n, k)).astype(np.float32)
    c = np.zeros((n, m), np.float32)

    def node(wt=3, kk=k, x_type=F32, dst_type=F32):
        s0 = tensor(lamm, wt, [kk, m], [lamm.type_size(wt) or 2, row_size(wt, kk) if wt != F16 else kk * 2], w)
        s1 = tensor(lamm, x_type, [kk, n], [4, kk * 4], x)
        d = tensor(lamm, dst_type, [m, n], [4, m * 4], c)
        d.src[0], d.src[1] = C.pointer(s0), C.pointer(s1)
        d._keep = (s0, s1)
        return d
    assert run(lamm, host, node())
    assert not run(lamm, host, node(), task=lamm.TASK_INIT)             # src/loongarch_matmul.cpp:15
    assert not run(lamm, host, node(), task=lamm.TASK_FINALIZE)
    assert not run(lamm, host, node(wt=F16, kk=k - 4))                  # F16 needs K % 8 == 0
    assert not run(lamm, host, node(wt=Q8_1))                           # an activation type is not a weight type
    assert not run(lamm, host, node(kk=k - 32 + 16))                    # K % QK
    assert not run(lamm, host, node(dst_type=F16))                      # :34
    assert not run(lamm, host, node(x_type=Q8_0))                       # Q4_1 pairs with Q8_1, not Q8_0


def test_only_leaf_weights_are_served_from_the_resident_copy(lamm, host, oracle):
    """ggml also hands the hook a MUTABLE src0 whose address and shape stay the same between calls: a view of the
    K/V cache, an activation or op result used as src0, a trainable (PARAM) or INPUT leaf.  The reference reads
    src0->data on every call (src/loongarch_matmul.cpp:131-135); here only a true leaf weight (op NONE, no
    view_src, no grad, neither INPUT nor PARAM) may hit the resident device copy -- everything else asks the
    registry for a refresh on every call."""
    rng = np.random.default_rng(5)
    m, n, k = 4, 2, 256
    w = oracle.quantize_weights(Q8_0, rng.standard_normal((m, k)).astype(np.float32))
    x = rng.standard_normal((n, k)).astype(np.float32)
    c = np.zeros((n, m), np.float32)
    parent = tensor(lamm, Q8_0, [k, m], [lamm.type_size(Q8_0), row_size(Q8_0, k)], w)

    def refreshes(**kw):
        s0 = tensor(lamm, Q8_0, [k, m], [lamm.type_size(Q8_0), row_size(Q8_0, k)], w)
        for key, v in kw.items():
            setattr(s0, key, v)
        s1 = tensor(lamm, F32, [k, n], [4, k * 4], x)
        d = tensor(lamm, F32, [m, n], [4, m * 4], c)
        d.src[0], d.src[1] = C.pointer(s0), C.pointer(s1)
        host.stub_reset()
        assert run(lamm, host, d)
        return host.stub_refreshes()
    assert refreshes() == 0                                             # a model weight
    assert refreshes(flags=2) == 0                                      # GGML_TENSOR_FLAG_OUTPUT alone changes nothing
    assert refreshes(view_src=C.pointer(parent)) == 1                   # K-cache view
    assert refreshes(op=24) == 1                                        # an op result as src0
    assert refreshes(flags=4) == 1                                      # GGML_TENSOR_FLAG_PARAM (finetune)
    assert refreshes(flags=1) == 1                                      # GGML_TENSOR_FLAG_INPUT
    assert refreshes(grad=C.pointer(parent)) == 1


@pytest.mark.parametrize("t", [3, Q8_0, Q2_K])
def test_prequantized_src1_and_wdata_writeback(lamm, host, oracle, t):
    """src1 already in the vec-dot type goes through lammcu_mul_mat_q_host; with the write-back hook on, params->wdata
    receives the quantized rows of every plane at ggml's packing (src/loongarch_matmul.cpp:137-138)."""
    rng = np.random.default_rng(40 + t)
    m, n, k, ne12 = 6, 3, 512, 2
    vt = VEC_DOT[t]
    rs, xrs = row_size(t, k), row_size(vt, k)
    w = oracle.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=2)
    x = rng.standard_normal((ne12, n, k)).astype(np.float32)
    xq = np.stack([oracle.quantize_act(vt, x[i]) for i in range(ne12)])
    c = np.zeros((ne12, n, m), np.float32)
    src0 = tensor(lamm, t, [k, m], [lamm.type_size(t), rs], w)
    src1q = tensor(lamm, vt, [k, n, ne12, 1], [lamm.type_size(vt) or {Q8_1: 36, Q8_K: 292}[vt], xrs, n * xrs], xq)
    dst = tensor(lamm, F32, [m, n, ne12, 1], [4, m * 4, n * m * 4], c)
    dst.src[0], dst.src[1] = C.pointer(src0), C.pointer(src1q)
    assert run(lamm, host, dst)
    for i in range(ne12):
        assert np.array_equal(c[i], oracle.mul_mat_q(t, w, xq[i], k))
    # f32 src1 + write-back
    src1 = tensor(lamm, F32, [k, n, ne12, 1], [4, k * 4, n * k * 4], x)
    dst.src[1] = C.pointer(src1)
    wdata = np.full((ne12 * n, xrs), 0, np.uint8)
    c[:] = 0
    host.lamm_set_writeback_wdata(1)
    try:
        assert run(lamm, host, dst, wdata=wdata)
    finally:
        host.lamm_set_writeback_wdata(0)
    assert np.array_equal(wdata.reshape(ne12, n, xrs), xq)
    for i in range(ne12):
        assert np.array_equal(c[i], oracle.mul_mat(t, w, x[i]))


@pytest.mark.parametrize("t", [3, Q8_0, Q2_K, 12, F32])
def test_mul_mat_id_grouping_gather_scatter(lamm, host, oracle, t):
    """lamm_mul_mat_id's host logic against orc_mul_mat_id (pinned to stock ggml in tests/test_oracle_vs_ref.py), bit for
    bit: one token; scattered tokens (gather + scatter, one call per expert that has tokens); tokens already consecutive
    per expert (no gather: the call reads src1 and writes dst in place); padded src1 / dst / ids strides; 2 threads."""
    rng = np.random.default_rng(70 + t)
    m, k, n_as, n_ids = 10, 512, 4, 3
    ws = [oracle.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=e).reshape(m, -1).view(np.uint8)
          if t != F32 else rng.standard_normal((m, k)).astype(np.float32).view(np.uint8).reshape(m, -1) for e in range(n_as)]
    rs = row_size(t, k)
    cases = [(1, np.array([[2, 0, 1]], np.int32)),
             (11, np.stack([rng.integers(0, n_as - 1, 11), np.full(11, 1), rng.integers(0, n_as, 11)], axis=1).astype(np.int32)),
             (9, np.stack([np.repeat([0, 1, 3], 3), np.repeat([3, 3, 0], 3), np.repeat([2, 2, 2], 3)], axis=1).astype(np.int32))]
    for n, ids in cases:
        ldx, ldc, ld_ids = k + 8, m + 3, n_ids + 2
        x = rng.standard_normal((n, ldx)).astype(np.float32)
        idp = np.full((n, ld_ids), -5, np.int32)
        idp[:, :n_ids] = ids
        for id_ in range(n_ids):
            c = np.full((n, ldc), 3.0, np.float32)
            tid = tensor(lamm, I32, [n_ids, n], [4, ld_ids * 4], idp)
            src1 = tensor(lamm, F32, [k, n], [4, ldx * 4], x)
            dst = tensor(lamm, F32, [m, n], [4, ldc * 4], c)
            dst.src[0], dst.src[1] = C.pointer(tid), C.pointer(src1)
            experts = [tensor(lamm, t, [k, m], [lamm.type_size(t), rs], w) for w in ws]
            for e, et in enumerate(experts):
                dst.src[2 + e] = C.pointer(et)
            dst.op_params[0], dst.op_params[1] = id_, n_as
            host.stub_reset()
            assert run(lamm, host, dst, nth=2, id_op=True)
            want = oracle.mul_mat_id(t, ws, ids, id_, x[:, :k])
            assert np.array_equal(c[:, :m], want), f"{TYPE_NAMES.get(t, t)} n {n} id {id_}"
            assert (c[:, m:] == 3.0).all()
            assert host.stub_calls() == len(set(ids[:, id_].tolist())) and host.stub_rows() == n
    # gate: wrong ids type, id outside the ids row, experts of different shapes, INIT phase
    n, ids = cases[1]
    x = rng.standard_normal((n, k)).astype(np.float32)
    c = np.zeros((n, m), np.float32)

    def node(ids_type=I32, id_=0, short_expert=False, n_as_=n_as):
        tid = tensor(lamm, ids_type, [n_ids, n], [4, n_ids * 4], ids)
        src1 = tensor(lamm, F32, [k, n], [4, k * 4], x)
        d = tensor(lamm, F32, [m, n], [4, m * 4], c)
        d.src[0], d.src[1] = C.pointer(tid), C.pointer(src1)
        ex = [tensor(lamm, t, [k, m - (1 if short_expert and e == 2 else 0)], [lamm.type_size(t), rs], w) for e, w in enumerate(ws)]
        for e, et in enumerate(ex):
            d.src[2 + e] = C.pointer(et)
        d.op_params[0], d.op_params[1] = id_, n_as_
        d._keep = (tid, src1, ex)
        return d
    assert run(lamm, host, node(), id_op=True)
    assert not run(lamm, host, node(ids_type=F32), id_op=True)
    assert not run(lamm, host, node(id_=n_ids), id_op=True)
    assert not run(lamm, host, node(short_expert=True), id_op=True)
    assert not run(lamm, host, node(n_as_=0), id_op=True)
    assert not run(lamm, host, node(n_as_=9), id_op=True)               # GGML_MAX_SRC - 2 = 8
    assert not run(lamm, host, node(), id_op=True, task=lamm.TASK_INIT)
