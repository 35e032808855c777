"""ctypes access to the CPU checkers (oracle/ and oracle/_ref).  Test infrastructure only.

`Oracle` wraps oracle/liblamm_oracle.so (our plain-C restatement, lamm_oracle.h).
`Ref` wraps oracle/_ref/libref.so (the unmodified reference compiled by
oracle/Makefile + ref_shim.cpp); it exists only where `make -C oracle ref` ran.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

# ggml type ids (llama.cpp-b2430/ggml.h:341-369)
F32, Q4_0, Q4_1, Q5_0, Q5_1, Q8_0, Q8_1, Q2_K, Q8_K = 0, 2, 3, 6, 7, 8, 9, 10, 15
Q3_K, Q4_K, Q5_K, Q6_K = 11, 12, 13, 14                   # SURVEY 8f-1 (dev.md:86-95)
TYPE_NAMES = {F32: "f32", Q4_0: "q4_0", Q4_1: "q4_1", Q5_0: "q5_0", Q5_1: "q5_1",
              Q8_0: "q8_0", Q8_1: "q8_1", Q2_K: "q2_k", Q8_K: "q8_k",
              Q3_K: "q3_k", Q4_K: "q4_k", Q5_K: "q5_k", Q6_K: "q6_k"}
NAME_TO_TYPE = {v: k for k, v in TYPE_NAMES.items()}
W_TYPES = [F32, Q2_K, Q4_0, Q4_1, Q5_0, Q5_1, Q8_0]     # src/loongarch_matmul.cpp:37-42
KQ_TYPES = [Q3_K, Q4_K, Q5_K, Q6_K]                     # beyond the reference's table: ggml's own vec_dot is the oracle
BLCK = {F32: 1, 1: 1, Q4_0: 32, Q4_1: 32, Q5_0: 32, Q5_1: 32, Q8_0: 32, Q8_1: 32, Q2_K: 256, Q8_K: 256,
        Q3_K: 256, Q4_K: 256, Q5_K: 256, Q6_K: 256}
TSIZE = {F32: 4, 1: 2, Q4_0: 18, Q4_1: 20, Q5_0: 22, Q5_1: 24, Q8_0: 34, Q8_1: 36, Q2_K: 84, Q8_K: 292,
         Q3_K: 110, Q4_K: 144, Q5_K: 176, Q6_K: 210}
VEC_DOT = {F32: F32, 1: 1, Q4_0: Q8_0, Q4_1: Q8_1, Q5_0: Q8_0, Q5_1: Q8_1, Q8_0: Q8_0, Q2_K: Q8_K,
           Q3_K: Q8_K, Q4_K: Q8_K, Q5_K: Q8_K, Q6_K: Q8_K}

# F16 weights x activations rounded to fp16 (ggml.c:486-496); in the product's table since round 2
F16 = 1
TYPE_NAMES[F16] = "f16"
NAME_TO_TYPE["f16"] = F16
ORACLE_ONLY = {F16: (1, 2, F16)}                          # type -> (block, bytes, vec-dot type), ggml.c:486-496


def vec_dot_of(t):
    return VEC_DOT[t] if t in VEC_DOT else ORACLE_ONLY[t][2]


i64, vp, fp = C.c_int64, C.c_void_p, C.POINTER(C.c_float)


def row_size(t, k):
    if t in ORACLE_ONLY:
        return (k // ORACLE_ONLY[t][0]) * ORACLE_ONLY[t][1]
    return (k // BLCK[t]) * TSIZE[t]


def _ptr(a):
    return a.ctypes.data_as(vp)


def _fptr(a):
    assert a.dtype == np.float32
    return a.ctypes.data_as(fp)


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "oracle"])


def build_ref():
    """Only possible where /root/reference exists (this container)."""
    if not os.path.isdir("/root/reference/src"):
        return False
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "ref"])
    return True


class Oracle:
    def __init__(self):
        path = os.path.join(ORACLE_DIR, "liblamm_oracle.so")
        if not os.path.exists(path):
            build_oracle()
        L = self.L = C.CDLL(path)
        L.orc_fp16_to_fp32.restype = C.c_float
        L.orc_fp16_to_fp32.argtypes = [C.c_uint16]
        L.orc_fp32_to_fp16.restype = C.c_uint16
        L.orc_fp32_to_fp16.argtypes = [C.c_float]
        L.orc_quantize_act.argtypes = [C.c_int, fp, i64, vp, i64, i64]
        L.orc_quantize_weights.argtypes = [C.c_int, fp, vp, i64, i64]
        L.orc_random_q2_K.argtypes = [vp, i64, C.c_uint64]
        L.orc_random_kquant.argtypes = [C.c_int, vp, i64, C.c_uint64]
        L.orc_dequantize_row.argtypes = [C.c_int, vp, fp, i64]
        L.orc_mul_mat_q.argtypes = [C.c_int, vp, i64, vp, i64, fp, i64, i64, i64, i64]
        L.orc_mul_mat_q_f32.argtypes = L.orc_mul_mat_q.argtypes
        L.orc_mul_mat.argtypes = [C.c_int, vp, i64, fp, i64, fp, i64, i64, i64, i64, vp]
        L.orc_mul_mat_mt.argtypes = [C.c_int, vp, i64, fp, i64, fp, i64, i64, i64, i64, C.c_int, vp]
        L.orc_mul_mat_id.argtypes = [C.c_int, C.POINTER(vp), C.c_int, i64, vp, i64, C.c_int, fp, i64, fp, i64,
                                     i64, i64, i64]

    def quantize_act(self, vt, x, prefill=0):
        """x: [N, K] f32 -> uint8 [N, row_size].  Output pre-filled (Q8_K bsums of
        all-zero blocks are never written)."""
        x = np.ascontiguousarray(x, np.float32)
        n, k = x.shape
        out = np.full((n, row_size(vt, k)), prefill, np.uint8)
        assert self.L.orc_quantize_act(vt, _fptr(x), k, _ptr(out), n, k) == 0
        return out

    def quantize_weights(self, t, w, seed=0):
        """w: [M, K] f32 -> uint8 [M, row_size].  Q2_K .. Q6_K: random super-blocks."""
        m, k = w.shape                       # Q2_K .. Q6_K: random super-blocks, only w.shape is used
        if t != Q2_K and t not in KQ_TYPES:
            w = np.ascontiguousarray(w, np.float32)
        out = np.zeros((m, row_size(t, k)), np.uint8)
        if t == Q2_K:
            self.L.orc_random_q2_K(_ptr(out), m * (k // 256), seed)
        elif t in KQ_TYPES:
            assert self.L.orc_random_kquant(t, _ptr(out), m * (k // 256), seed) == 0
        else:
            assert self.L.orc_quantize_weights(t, _fptr(w), _ptr(out), m, k) == 0
        return out

    def dequantize(self, t, q, k):
        q = np.ascontiguousarray(q)
        rows = q.reshape(-1, row_size(t, k))
        out = np.zeros((rows.shape[0], k), np.float32)
        for r in range(rows.shape[0]):
            assert self.L.orc_dequantize_row(t, _ptr(rows[r]), _fptr(out[r]), k) == 0
        return out

    def mul_mat_q(self, t, wq, xq, k, f32_order=False):
        """wq: uint8 [M, row_size(t)], xq: uint8 [N, row_size(vec_dot)] -> C [N, M] f32"""
        wq = np.ascontiguousarray(wq)
        xq = np.ascontiguousarray(xq)
        m, n = wq.shape[0], xq.shape[0]
        c = np.zeros((n, m), np.float32)
        fn = self.L.orc_mul_mat_q_f32 if f32_order else self.L.orc_mul_mat_q
        assert fn(t, _ptr(wq), wq.strides[0], _ptr(xq), xq.strides[0], _fptr(c), m, m, n, k) == 0
        return c

    def mul_mat(self, t, wq, x, want_xq=False):
        wq = np.ascontiguousarray(wq)
        x = np.ascontiguousarray(x, np.float32)
        m, (n, k) = wq.shape[0], x.shape
        c = np.zeros((n, m), np.float32)
        xq = np.zeros((n, row_size(vec_dot_of(t), k)), np.uint8)
        assert self.L.orc_mul_mat(t, _ptr(wq), wq.strides[0], _fptr(x), k, _fptr(c), m, m, n, k, _ptr(xq)) == 0
        return (c, xq) if want_xq else c

    def mul_mat_id(self, t, wqs, ids, id_, x):
        """wqs: list of uint8 [M, row_size] expert matrices, ids: int32 [N, n_ids], x: [N, K] -> C [N, M]"""
        wqs = [np.ascontiguousarray(w) for w in wqs]
        ids = np.ascontiguousarray(ids, np.int32)
        x = np.ascontiguousarray(x, np.float32)
        m, (n, k) = wqs[0].shape[0], x.shape
        c = np.zeros((n, m), np.float32)
        ptrs = (vp * len(wqs))(*[w.ctypes.data for w in wqs])
        rc = self.L.orc_mul_mat_id(t, ptrs, len(wqs), wqs[0].strides[0], _ptr(ids), ids.shape[1], id_, _fptr(x), k,
                                   _fptr(c), m, m, n, k)
        assert rc == 0, rc
        return c

    def rms_norm_mul(self, x, norm_w, eps):
        """ggml_rms_norm then ggml_mul by the norm weight (None: no weight) -> f32 [N, K]"""
        x = np.ascontiguousarray(x, np.float32)
        n, k = x.shape
        y = np.zeros((n, k), np.float32)
        w = None if norm_w is None else np.ascontiguousarray(norm_w, np.float32)
        self.L.orc_rms_norm_mul.argtypes = [fp, i64, fp, C.c_float, fp, i64, i64, i64]
        assert self.L.orc_rms_norm_mul(_fptr(x), k, None if w is None else _fptr(w), eps, _fptr(y), k, n, k) == 0
        return y

    def silu_mul(self, gate, up):
        """ggml_silu(gate) * up (fp16 table SiLU, ggml.c:1894-1902) -> f32 [N, K]"""
        gate = np.ascontiguousarray(gate, np.float32)
        up = np.ascontiguousarray(up, np.float32)
        n, k = gate.shape
        y = np.zeros((n, k), np.float32)
        self.L.orc_silu_mul.argtypes = [fp, i64, fp, i64, fp, i64, i64, i64]
        assert self.L.orc_silu_mul(_fptr(gate), k, _fptr(up), k, _fptr(y), k, n, k) == 0
        return y

    def mul_mat_mt(self, t, wq, x, nth):
        wq = np.ascontiguousarray(wq)
        x = np.ascontiguousarray(x, np.float32)
        m, (n, k) = wq.shape[0], x.shape
        c = np.zeros((n, m), np.float32)
        assert self.L.orc_mul_mat_mt(t, _ptr(wq), wq.strides[0], _fptr(x), k, _fptr(c), m, m, n, k, nth, None) == 0
        return c


class Ref:
    """The real reference (oracle/_ref/libref.so)."""
    LIBNAME = "libref.so"

    @classmethod
    def available(cls):
        return os.path.exists(os.path.join(ORACLE_DIR, "_ref", cls.LIBNAME))

    def __init__(self):
        L = self.L = C.CDLL(os.path.join(ORACLE_DIR, "_ref", self.LIBNAME))
        L.ref_init()
        if hasattr(L, "ref_graph_mul_mat_wdata"):
            L.ref_graph_mul_mat_wdata.argtypes = [C.c_int, vp, fp, fp, vp, C.c_size_t, i64, i64, i64, C.c_int, C.c_int, C.c_int]
        L.ref_quantize_act.argtypes = [C.c_int, fp, i64, vp, i64, i64]
        L.ref_quantize_act_native.argtypes = L.ref_quantize_act.argtypes
        L.ref_quantize_weights.argtypes = [C.c_int, fp, vp, i64, i64]
        L.ref_dequantize_row.argtypes = [C.c_int, vp, fp, i64]
        L.ref_vec_dot.restype = C.c_float
        L.ref_vec_dot.argtypes = [C.c_int, vp, vp, i64]
        L.ref_lamm_mul_mat.argtypes = [C.c_int, vp, i64, fp, vp, fp, i64, i64, i64, i64, C.c_int]
        L.ref_graph_mul_mat.argtypes = [C.c_int, vp, fp, fp, i64, i64, i64, C.c_int]
        if hasattr(L, "ref_graph_mul_mat_id"):
            L.ref_graph_mul_mat_id.argtypes = [C.c_int, vp, C.c_int, vp, C.c_int, C.c_int, fp, fp, i64, i64, i64, C.c_int]
        L.ref_bench_mul_mat.argtypes = [C.c_int, vp, fp, i64, i64, i64, C.c_int, C.c_int, C.POINTER(C.c_double)]
        L.ref_row_size.restype = C.c_size_t
        L.ref_row_size.argtypes = [C.c_int, i64]

    def quantize_act(self, vt, x, prefill=0, native=False):
        x = np.ascontiguousarray(x, np.float32)
        n, k = x.shape
        out = np.full((n, row_size(vt, k)), prefill, np.uint8)
        fn = self.L.ref_quantize_act_native if native else self.L.ref_quantize_act
        assert fn(vt, _fptr(x), k, _ptr(out), n, k) == 0
        return out

    def quantize_weights(self, t, w):
        w = np.ascontiguousarray(w, np.float32)
        m, k = w.shape
        out = np.zeros((m, row_size(t, k)), np.uint8)
        assert self.L.ref_quantize_weights(t, _fptr(w), _ptr(out), m, k) == 0
        return out

    def dequantize(self, t, q, k):
        rows = np.ascontiguousarray(q).reshape(-1, row_size(t, k))
        out = np.zeros((rows.shape[0], k), np.float32)
        for r in range(rows.shape[0]):
            assert self.L.ref_dequantize_row(t, _ptr(rows[r]), _fptr(out[r]), k) == 0
        return out

    def vec_dot(self, t, wq, xq, k):
        wq = np.ascontiguousarray(wq)
        xq = np.ascontiguousarray(xq)
        c = np.zeros((xq.shape[0], wq.shape[0]), np.float32)
        for j in range(xq.shape[0]):
            for i in range(wq.shape[0]):
                c[j, i] = self.L.ref_vec_dot(t, _ptr(wq[i]), _ptr(xq[j]), k)
        return c

    def lamm_mul_mat(self, t, wq, x, xq, nth=1):
        """The reference operator called directly on pre-quantized activations."""
        wq = np.ascontiguousarray(wq)
        x = np.ascontiguousarray(x, np.float32)
        xq = np.ascontiguousarray(xq)
        m, (n, k) = wq.shape[0], x.shape
        c = np.zeros((n, m), np.float32)
        rc = self.L.ref_lamm_mul_mat(t, _ptr(wq), wq.strides[0], _fptr(x), _ptr(xq), _fptr(c), m, m, n, k, nth)
        assert rc == 0, "lamm_can_mul_mat refused"
        return c

    def graph_mul_mat(self, t, wq, x, nth=1):
        """The reference's whole path through ggml_graph_compute."""
        wq = np.ascontiguousarray(wq)
        x = np.ascontiguousarray(x, np.float32)
        m, (n, k) = wq.shape[0], x.shape
        c = np.zeros((n, m), np.float32)
        assert self.L.ref_graph_mul_mat(t, _ptr(wq), _fptr(x), _fptr(c), m, n, k, nth) == 0
        return c

    def graph_mul_mat_wdata(self, t, wq, x, nth=1, reps=1, writeback=0):
        """As graph_mul_mat; also returns the first N * row_size(vec_dot, K) bytes of the plan's work buffer
        (params->wdata: the quantized activations) as they are after the last compute."""
        wq = np.ascontiguousarray(wq)
        x = np.ascontiguousarray(x, np.float32)
        m, (n, k) = wq.shape[0], x.shape
        c = np.zeros((n, m), np.float32)
        wd = np.zeros((n, row_size(vec_dot_of(t), k) if t != F32 else 4), np.uint8)
        assert self.L.ref_graph_mul_mat_wdata(t, _ptr(wq), _fptr(x), _fptr(c), _ptr(wd), wd.nbytes, m, n, k, nth, reps,
                                              writeback) == 0
        return c, wd

    def graph_mul_mat_id(self, t, wqs, ids, id_, x, nth=1):
        """ggml_mul_mat_id through ggml_graph_compute (stock ggml; lamm's hook is not on this op)."""
        ws = np.ascontiguousarray(np.stack([np.ascontiguousarray(w) for w in wqs]))
        ids = np.ascontiguousarray(ids, np.int32)
        x = np.ascontiguousarray(x, np.float32)
        m, (n, k) = wqs[0].shape[0], x.shape
        c = np.zeros((n, m), np.float32)
        rc = self.L.ref_graph_mul_mat_id(t, _ptr(ws), len(wqs), _ptr(ids), ids.shape[1], id_, _fptr(x), _fptr(c),
                                         m, n, k, nth)
        assert rc == 0, rc
        return c

    def graph_xop_mul_mat(self, kind, t, wq, x, aux, eps=0.0, nth=1):
        """SURVEY 8f-3 through stock ggml's graph: kind 1 = rms_norm(x, eps) * aux[K] (aux None: no weight), kind 2 =
        silu(x) * aux[N, K]; then ggml_mul_mat(w, .).  Returns (C [N, M], y [N, K] = what the matmul consumed)."""
        wq = np.ascontiguousarray(wq)
        x = np.ascontiguousarray(x, np.float32)
        m, (n, k) = wq.shape[0], x.shape
        c = np.zeros((n, m), np.float32)
        y = np.zeros((n, k), np.float32)
        a = None if aux is None else np.ascontiguousarray(aux, np.float32)
        self.L.ref_graph_xop_mul_mat.argtypes = [C.c_int, C.c_int, vp, fp, fp, C.c_float, fp, fp, i64, i64, i64, C.c_int]
        rc = self.L.ref_graph_xop_mul_mat(kind, t, _ptr(wq), _fptr(x), None if a is None else _fptr(a), eps, _fptr(c),
                                          _fptr(y), m, n, k, nth)
        assert rc == 0, rc
        return c, y

    def gguf_write(self, path, tensors, alignment=0):
        """tensors: list of (name, ggml type, numpy array [ne1, row bytes] or 1-D bytes, ne0, ne1 (0 = 1-D)) written with
        the reference's own gguf writer, preceded by a fixed set of key/values of every kind."""
        n = len(tensors)
        names = (C.c_char_p * n)(*[t[0].encode() for t in tensors])
        types = (C.c_int * n)(*[t[1] for t in tensors])
        arrs = [np.ascontiguousarray(t[2]) for t in tensors]
        ne = (i64 * (2 * n))(*[v for t in tensors for v in (t[3], t[4])])
        data = (vp * n)(*[a.ctypes.data for a in arrs])
        self.L.ref_gguf_write.argtypes = [C.c_char_p, C.c_uint32, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int),
                                          C.POINTER(i64), C.POINTER(vp)]
        assert self.L.ref_gguf_write(path.encode(), alignment, n, names, types, ne, data) == 0

    def bench(self, t, m, n, k, nth, iters, wq=None, x=None):
        us = (C.c_double * iters)()
        wp = _ptr(np.ascontiguousarray(wq)) if wq is not None else None
        xp = _fptr(np.ascontiguousarray(x, np.float32)) if x is not None else None
        assert self.L.ref_bench_mul_mat(t, wp, xp, m, n, k, nth, iters, us) == 0
        return list(us)


class RefB200(Ref):
    """oracle/_ref/libref_b200.so: the reference's own ggml.c (-DLA_LLAMA) + ref_shim.cpp linked against
    la-llama.cpp_b200/liblammcu.so instead of the reference's loongarch_matmul.o (oracle/Makefile, ref_b200).
    ggml_graph_compute's real thread pool calls the B200 operator through the hook at ggml.c:10858-10863."""
    LIBNAME = "libref_b200.so"

    def __init__(self):
        super().__init__()
        assert self.L.ref_is_b200() == 1, "libref_b200.so is not linked against liblammcu.so"
