"""la-llama.cpp_b200 -- B200 (sm_100a) quantized mul_mat behind the LAMM operator boundary.

Python view of the C-ABI in include/lammcu.h and of the drop-in operator in
include/lamm_b200.h (`lamm_can_mul_mat` / `lamm_mul_mat` / `lamm_get_opt_level`,
reference: src/loongarch_matmul.h:18-24).  This module is plumbing only: it loads
liblammcu.so with ctypes and passes raw pointers; all arithmetic runs in the CUDA
library.  There is no CPU fallback -- if the library is missing or no B200 is
visible, compute calls raise `LammError`.

The directory name is not a valid Python identifier; import it through the
repo-root shim `lamm_b200.py` (``import lamm_b200``).
"""
import ctypes as C
import os

from . import build as _build

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LAMM_LIB") or os.path.join(HERE, "liblammcu.so")     # LAMM_LIB: a variant build (kernel experiments)

# ggml type ids (llama.cpp-b2430/ggml.h:341-369)
F32, Q4_0, Q4_1, Q5_0, Q5_1, Q8_0, Q8_1, Q2_K, Q8_K = 0, 2, 3, 6, 7, 8, 9, 10, 15
Q3_K, Q4_K, Q5_K, Q6_K = 11, 12, 13, 14                   # beyond the reference's table
TYPE_NAMES = {F32: "f32", Q4_0: "q4_0", Q4_1: "q4_1", Q5_0: "q5_0", Q5_1: "q5_1",
              Q8_0: "q8_0", Q8_1: "q8_1", Q2_K: "q2_k", Q8_K: "q8_k",
              Q3_K: "q3_k", Q4_K: "q4_k", Q5_K: "q5_k", Q6_K: "q6_k"}
NAME_TO_TYPE = {v: k for k, v in TYPE_NAMES.items()}
W_TYPES = (F32, Q2_K, Q4_0, Q4_1, Q5_0, Q5_1, Q8_0)      # src/loongarch_matmul.cpp:37-42
KQ_TYPES = (Q3_K, Q4_K, Q5_K, Q6_K)
F16 = 1                                                   # F16 weights, activations rounded to fp16 (ggml.c:486-496)
TYPE_NAMES[F16] = "f16"
NAME_TO_TYPE["f16"] = F16
TASK_INIT, TASK_COMPUTE, TASK_FINALIZE = 0, 1, 2

# every symbol include/lammcu.h and include/lamm_b200.h declare
EXPORTS = [
    "lammcu_version", "lammcu_last_error", "lammcu_device_count", "lammcu_init", "lammcu_shutdown",
    "lammcu_sm_count", "lammcu_launch_count", "lammcu_launch_count_reset", "lammcu_type_supported",
    "lammcu_blck_size", "lammcu_type_size", "lammcu_vec_dot_type", "lammcu_row_size",
    "lammcu_quantize_act_dev", "lammcu_mul_mat_q_dev", "lammcu_mul_mat_dev", "lammcu_set_gemv_max_n",
    "lammcu_get_gemv_max_n", "lammcu_set_decode_kernel", "lammcu_get_decode_kernel", "lammcu_set_trace", "lammcu_ipc_alloc", "lammcu_ipc_open", "lammcu_ipc_close", "lammcu_ipc_free", "lammcu_memcpy_async",
    "lammcu_mul_mat_batch_dev", "lammcu_quantize_act_xop_dev", "lammcu_mul_mat_xop_dev", "lammcu_mul_mat_batch_xop_dev", "lammcu_mul_mat_noalloc_dev", "lammcu_chain_create", "lammcu_chain_run_dev", "lammcu_chain_status", "lammcu_chain_destroy", "lammcu_set_gemm_exact", "lammcu_get_gemm_exact", "lammcu_host_profile", "lammcu_set_host_mode", "lammcu_get_host_mode", "lammcu_host_quiesce", "lammcu_set_server_timeouts", "lammcu_server_trace", "lammcu_comm_create", "lammcu_comm_destroy", "lammcu_comm_status", "lammcu_comm_sync", "lammcu_ll_bytes", "lammcu_mul_mat_sharded_dev", "lammcu_mul_mat_sharded_batch_dev", "lammcu_mul_mat_sharded_gemm_dev", "lammcu_ll_unpack_dev", "lammcu_weight_register", "lammcu_weight_for_op", "lammcu_weight_lookup", "lammcu_weight_release", "lammcu_weight_forget", "lammcu_weight_forget_all",
    "lammcu_weight_dev_ptr", "lammcu_weight_rows", "lammcu_weight_bytes", "lammcu_mul_mat_host", "lammcu_mul_mat_q_host",
    "lamm_can_mul_mat", "lamm_mul_mat", "lamm_get_opt_level", "lamm_set_writeback_wdata",
    "lammcu_mul_mat_id_plan", "lamm_can_mul_mat_id", "lamm_mul_mat_id",
    "lammcu_gguf_open", "lammcu_gguf_close", "lammcu_gguf_tensor_count", "lammcu_gguf_alignment",
    "lammcu_gguf_data_offset", "lammcu_gguf_tensor", "lammcu_gguf_find", "lammcu_gguf_get_u32", "lammcu_gguf_get_f32",
    "lammcu_gguf_get_str", "lammcu_gguf_upload_rows", "lammcu_gguf_rows",
]


class LammError(RuntimeError):
    pass


# ---- ggml structs (include/lamm_b200.h; llama.cpp-b2430/ggml.h:552-590, 668-677) ----
class GgmlComputeParams(C.Structure):
    _fields_ = [("type", C.c_int), ("ith", C.c_int), ("nth", C.c_int),
                ("wsize", C.c_size_t), ("wdata", C.c_void_p)]


class GgmlTensor(C.Structure):
    pass


GgmlTensor._fields_ = [
    ("type", C.c_int), ("backend", C.c_int), ("buffer", C.c_void_p),
    ("ne", C.c_int64 * 4), ("nb", C.c_size_t * 4), ("op", C.c_int),
    ("op_params", C.c_int32 * 16), ("flags", C.c_int32),
    ("grad", C.POINTER(GgmlTensor)), ("src", C.POINTER(GgmlTensor) * 10),
    ("perf_runs", C.c_int), ("perf_cycles", C.c_int64), ("perf_time_us", C.c_int64),
    ("view_src", C.POINTER(GgmlTensor)), ("view_offs", C.c_size_t),
    ("data", C.c_void_p), ("name", C.c_char * 64), ("extra", C.c_void_p), ("padding", C.c_char * 8),
]

class ChainOp(C.Structure):
    """lammcu_chain_op (include/lammcu.h)"""
    _fields_ = [("count", C.c_int), ("w_dev", C.c_void_p * 4), ("M", C.c_int64 * 4), ("c_dev", C.c_void_p * 4),
                ("x_dev", C.c_void_p), ("K", C.c_int64), ("w_stride_bytes", C.c_int64)]


_lib = None


def lib():
    """Load liblammcu.so (building it if the sources are newer).  Raises LammError."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        try:
            _build.build()
        except Exception as e:  # no nvcc, compile error
            raise LammError(f"liblammcu.so is missing and could not be built: {e}") from e
    try:
        L = C.CDLL(LIB_PATH)
    except OSError as e:
        raise LammError(f"cannot load {LIB_PATH}: {e}") from e
    i64, vp, fp, sz = C.c_int64, C.c_void_p, C.c_void_p, C.c_size_t
    L.lammcu_version.restype = C.c_char_p
    L.lammcu_last_error.restype = C.c_char_p
    L.lammcu_init.argtypes = [C.c_int]
    L.lammcu_launch_count.restype = C.c_uint64
    L.lammcu_launch_count_reset.restype = None
    L.lammcu_type_size.restype = sz
    L.lammcu_type_size.argtypes = [C.c_int]
    L.lammcu_blck_size.argtypes = [C.c_int]
    L.lammcu_vec_dot_type.argtypes = [C.c_int]
    L.lammcu_type_supported.argtypes = [C.c_int]
    L.lammcu_row_size.restype = sz
    L.lammcu_row_size.argtypes = [C.c_int, i64]
    L.lammcu_quantize_act_dev.argtypes = [C.c_int, fp, i64, vp, i64, i64, vp]
    L.lammcu_mul_mat_q_dev.argtypes = [C.c_int, vp, i64, vp, i64, fp, i64, i64, i64, i64, vp]
    L.lammcu_mul_mat_dev.argtypes = [C.c_int, vp, i64, fp, i64, fp, i64, i64, i64, i64, vp, vp]
    L.lammcu_mul_mat_noalloc_dev.argtypes = [C.c_int, vp, i64, fp, i64, fp, i64, i64, i64, i64, vp]
    L.lammcu_chain_create.argtypes = [C.c_int, C.c_int, C.POINTER(ChainOp), C.POINTER(vp)]
    L.lammcu_chain_run_dev.argtypes = [vp, vp]
    L.lammcu_chain_status.argtypes = [vp, C.POINTER(C.c_int)]
    L.lammcu_chain_destroy.argtypes = [vp]
    L.lammcu_set_gemv_max_n.argtypes = [C.c_int]
    L.lammcu_set_decode_kernel.argtypes = [C.c_int]
    L.lammcu_set_trace.argtypes = [vp, i64]
    L.lammcu_ipc_alloc.argtypes = [sz, C.POINTER(vp), vp]
    L.lammcu_ipc_open.argtypes = [vp, C.POINTER(vp)]
    L.lammcu_ipc_close.argtypes = [vp]
    L.lammcu_ipc_free.argtypes = [vp]
    L.lammcu_memcpy_async.argtypes = [vp, vp, sz, vp]
    L.lammcu_mul_mat_batch_dev.argtypes = [C.c_int, C.c_int, C.POINTER(vp), i64, C.POINTER(i64), fp, C.POINTER(vp), i64, vp, vp]
    L.lammcu_quantize_act_xop_dev.argtypes = [vp, C.c_int, vp, i64, vp, i64, i64, vp]
    L.lammcu_mul_mat_xop_dev.argtypes = [vp, C.c_int, vp, i64, vp, i64, vp, i64, i64, i64, i64, vp, vp]
    L.lammcu_mul_mat_batch_xop_dev.argtypes = [vp, C.c_int, C.c_int, C.POINTER(vp), i64, C.POINTER(i64), vp, C.POINTER(vp), i64, vp, vp]
    L.lammcu_host_profile.restype = None
    L.lammcu_host_profile.argtypes = [C.POINTER(C.c_double), C.c_int]
    L.lammcu_comm_create.argtypes = [C.c_int, C.c_int, C.POINTER(vp)]
    L.lammcu_comm_destroy.argtypes = [vp]
    L.lammcu_comm_status.argtypes = [vp, C.POINTER(C.c_int)]
    L.lammcu_comm_sync.argtypes = [vp, vp]
    L.lammcu_ll_bytes.restype = C.c_size_t
    L.lammcu_ll_bytes.argtypes = [i64]
    L.lammcu_mul_mat_sharded_dev.argtypes = [vp, C.c_int, vp, i64, vp, C.c_int, C.POINTER(vp), i64, i64, i64, vp]
    L.lammcu_mul_mat_sharded_batch_dev.argtypes = [vp, C.c_int, C.c_int, C.POINTER(vp), i64, vp, C.c_int, C.POINTER(vp),
                                                   C.POINTER(i64), C.POINTER(i64), i64, vp]
    L.lammcu_mul_mat_sharded_gemm_dev.argtypes = [vp, C.c_int, vp, i64, fp, i64, C.POINTER(vp), i64, C.POINTER(vp),
                                                  i64, i64, i64, i64, vp]
    L.lammcu_ll_unpack_dev.argtypes = [vp, vp, C.c_int, fp, i64, vp]
    L.lammcu_weight_register.argtypes = [vp, C.c_int, i64, i64, i64, i64, i64, C.POINTER(vp)]
    L.lammcu_weight_for_op.argtypes = [vp, C.c_int, i64, i64, i64, C.c_int, C.POINTER(vp)]
    L.lammcu_weight_lookup.argtypes = [vp, C.POINTER(vp)]
    L.lammcu_weight_release.argtypes = [vp]
    L.lammcu_weight_forget.argtypes = [vp]
    L.lammcu_weight_dev_ptr.restype = vp
    L.lammcu_weight_dev_ptr.argtypes = [vp]
    L.lammcu_weight_rows.restype = i64
    L.lammcu_weight_rows.argtypes = [vp]
    L.lammcu_weight_bytes.restype = i64
    L.lammcu_weight_bytes.argtypes = [vp]
    L.lammcu_mul_mat_host.argtypes = [vp, fp, i64, i64, fp, i64, vp]
    L.lammcu_mul_mat_q_host.argtypes = [vp, vp, i64, i64, fp, i64]
    L.lamm_can_mul_mat.restype = C.c_bool
    L.lamm_can_mul_mat.argtypes = [C.POINTER(GgmlComputeParams), C.POINTER(GgmlTensor)]
    L.lamm_mul_mat.restype = None
    L.lamm_mul_mat.argtypes = [C.POINTER(GgmlComputeParams), C.POINTER(GgmlTensor)]
    L.lamm_can_mul_mat_id.restype = C.c_bool
    L.lamm_can_mul_mat_id.argtypes = [C.POINTER(GgmlComputeParams), C.POINTER(GgmlTensor)]
    L.lamm_mul_mat_id.restype = None
    L.lamm_mul_mat_id.argtypes = [C.POINTER(GgmlComputeParams), C.POINTER(GgmlTensor)]
    L.lammcu_gguf_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    L.lammcu_gguf_close.argtypes = [C.c_void_p]
    L.lammcu_gguf_tensor_count.restype = C.c_int64
    L.lammcu_gguf_tensor_count.argtypes = [C.c_void_p]
    L.lammcu_gguf_alignment.restype = C.c_uint32
    L.lammcu_gguf_alignment.argtypes = [C.c_void_p]
    L.lammcu_gguf_data_offset.restype = C.c_uint64
    L.lammcu_gguf_data_offset.argtypes = [C.c_void_p]
    L.lammcu_gguf_tensor.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_char_p), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                     C.POINTER(C.c_int64), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    L.lammcu_gguf_find.restype = C.c_int64
    L.lammcu_gguf_find.argtypes = [C.c_void_p, C.c_char_p]
    L.lammcu_gguf_get_u32.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_uint32)]
    L.lammcu_gguf_get_f32.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_float)]
    L.lammcu_gguf_get_str.restype = C.c_int64
    L.lammcu_gguf_get_str.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int64]
    L.lammcu_gguf_rows.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    L.lammcu_gguf_upload_rows.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.POINTER(C.c_void_p)]
    L.lammcu_mul_mat_id_plan.restype = C.c_int
    L.lammcu_mul_mat_id_plan.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]
    L.lamm_set_writeback_wdata.restype = None
    L.lamm_set_writeback_wdata.argtypes = [C.c_int]
    _lib = L
    return L


def check(rc, what=""):
    if rc != 0:
        raise LammError(f"{what or 'lammcu'} failed ({rc}): {lib().lammcu_last_error().decode()}")


def type_size(t):
    return lib().lammcu_type_size(t)


def blck_size(t):
    return lib().lammcu_blck_size(t)


def vec_dot_type(t):
    return lib().lammcu_vec_dot_type(t)


def row_size(t, k):
    return lib().lammcu_row_size(t, k)


def init(device=-1):
    check(lib().lammcu_init(device), "lammcu_init")


# ---- device-tensor helpers (torch = device memory + streams only) -----------------
CUDA_STREAM_LEGACY = 1      # cudaStreamLegacy: the NULL stream, spelled so it is not "library's own"


def _stream_ptr():
    """torch's current stream as the void* the C-ABI wants.  torch's default stream has handle 0,
    which lammcu.h reads as 'use the library's own stream'; pass cudaStreamLegacy instead so the
    kernels stay ordered with torch's copies on that stream."""
    import torch
    h = torch.cuda.current_stream().cuda_stream
    return C.c_void_p(h if h else CUDA_STREAM_LEGACY)


def quantize_act(x, vt):
    """x: cuda f32 [N, K] (row stride in elements = x.stride(0)) -> cuda uint8 [N, row_size]."""
    import torch
    assert x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
    n, k = x.shape
    out = torch.empty((n, row_size(vt, k)), dtype=torch.uint8, device=x.device)
    check(lib().lammcu_quantize_act_dev(vt, x.data_ptr(), x.stride(0), out.data_ptr(), n, k, _stream_ptr()),
          "quantize_act")
    return out


def mul_mat_q(w_type, w, xq, k, out=None):
    """w: cuda uint8 [M, row bytes] (F32: float32 [M, K]); xq: cuda uint8 [N, row_size(vec_dot)]
    (F32: float32 [N, K]).  Returns C: cuda f32 [N, M]."""
    import torch
    m, n = w.shape[0], xq.shape[0]
    if out is None:
        out = torch.empty((n, m), dtype=torch.float32, device=w.device)
    check(lib().lammcu_mul_mat_q_dev(w_type, w.data_ptr(), w.stride(0) * w.element_size(), xq.data_ptr(),
                                     xq.stride(0) * xq.element_size(), out.data_ptr(), out.stride(0),
                                     m, n, k, _stream_ptr()), "mul_mat_q")
    return out


def mul_mat(w_type, w, x, out=None, ws=None):
    """The whole path on device tensors: fused activation quantize + mul_mat.
    w: cuda uint8 [M, row bytes] in ggml block layout; x: cuda f32 [N, K]; returns f32 [N, M]."""
    import torch
    assert x.is_cuda and x.dtype == torch.float32 and x.stride(1) == 1
    m, (n, k) = w.shape[0], x.shape
    if out is None:
        out = torch.empty((n, m), dtype=torch.float32, device=x.device)
    check(lib().lammcu_mul_mat_dev(w_type, w.data_ptr(), w.stride(0) * w.element_size(), x.data_ptr(),
                                   x.stride(0), out.data_ptr(), out.stride(0), m, n, k,
                                   ws.data_ptr() if ws is not None else None, _stream_ptr()), "mul_mat")
    return out


def mul_mat_batch(w_type, ws, x, outs):
    """Sibling matmuls sharing one activation row (N = 1): outs[j] = ws[j] . x, one launch when possible.
    ws: cuda uint8 [M_j, row bytes] with equal row stride; x: cuda f32 [1, K]; outs: cuda f32 [1, M_j]."""
    n = len(ws)
    k = x.shape[-1]
    stride = ws[0].stride(0) * ws[0].element_size()
    assert all(w.stride(0) * w.element_size() == stride for w in ws)
    wp = (C.c_void_p * n)(*[w.data_ptr() for w in ws])
    cp = (C.c_void_p * n)(*[o.data_ptr() for o in outs])
    ms = (C.c_int64 * n)(*[w.shape[0] for w in ws])
    check(lib().lammcu_mul_mat_batch_dev(w_type, n, wp, stride, ms, x.data_ptr(), cp, k, None, _stream_ptr()),
          "mul_mat_batch")
    return outs


# ---- the op in front of the matmul, fused into the activation quantizer ----
XOP_RMS_NORM_MUL, XOP_SILU_MUL = 1, 2


class XOp(C.Structure):
    """include/lammcu.h: lammcu_xop"""
    _fields_ = [("kind", C.c_int), ("aux", C.c_void_p), ("ld_aux", C.c_int64), ("eps", C.c_float)]


def rms_norm_op(norm_w, eps):
    """llm_build_norm: ggml_rms_norm(x, eps) * norm_w (cuda f32 [K], or None)"""
    op = XOp(XOP_RMS_NORM_MUL, norm_w.data_ptr() if norm_w is not None else None, 0, eps)
    op._keep = norm_w
    return op


def silu_mul_op(up):
    """llm_build_ffn: silu(x) * up (cuda f32 [N, K], row stride up.stride(0))"""
    op = XOp(XOP_SILU_MUL, up.data_ptr(), up.stride(0), 0.0)
    op._keep = up
    return op


def quantize_act_xop(op, x, vt):
    """quantize(vt, op(x)): x cuda f32 [N, K] -> cuda uint8 [N, row_size(vt, K)] (vt F32: the f32 bytes of op(x))"""
    import torch
    n, k = x.shape
    rs = k * 4 if vt == F32 else row_size(vt, k)
    out = torch.empty((n, rs), dtype=torch.uint8, device=x.device)
    check(lib().lammcu_quantize_act_xop_dev(C.byref(op), vt, x.data_ptr(), x.stride(0), out.data_ptr(), n, k,
                                            _stream_ptr()), "quantize_act_xop")
    return out


def mul_mat_xop(op, w_type, w, x, out=None, ws=None):
    """C = W . op(x): mul_mat with the op fused in front (N == 1: one launch)."""
    import torch
    m, (n, k) = w.shape[0], x.shape
    if out is None:
        out = torch.empty((n, m), dtype=torch.float32, device=x.device)
    check(lib().lammcu_mul_mat_xop_dev(C.byref(op), w_type, w.data_ptr(), w.stride(0) * w.element_size(), x.data_ptr(),
                                       x.stride(0), out.data_ptr(), out.stride(0), m, n, k,
                                       ws.data_ptr() if ws is not None else None, _stream_ptr()), "mul_mat_xop")
    return out


def mul_mat_batch_xop(op, w_type, ws, x, outs):
    """Sibling matmuls sharing op(x) (N = 1): outs[j] = ws[j] . op(x), one launch when possible."""
    n = len(ws)
    k = x.shape[-1]
    stride = ws[0].stride(0) * ws[0].element_size()
    wp = (C.c_void_p * n)(*[w.data_ptr() for w in ws])
    cp = (C.c_void_p * n)(*[o.data_ptr() for o in outs])
    ms = (C.c_int64 * n)(*[w.shape[0] for w in ws])
    check(lib().lammcu_mul_mat_batch_xop_dev(C.byref(op), w_type, n, wp, stride, ms, x.data_ptr(), cp, k, None,
                                             _stream_ptr()), "mul_mat_batch_xop")
    return outs


# ---- the drop-in operator on host (numpy) tensors, as ggml would call it -------------
class Chain:
    """lammcu_chain_*: a list of dependent N == 1 mul_mats as one persistent launch.  ops: list of
    (weights, x, outs): weights = list of 1..4 uint8 [M_j, row bytes] device tensors of one type / K, x = f32 [1, K]
    device tensor, outs = list of f32 [1, M_j] device tensors.  The tensors must outlive the chain."""

    def __init__(self, w_type, ops):
        L = lib()
        self._keep = ops
        arr = (ChainOp * len(ops))()
        for i, (ws, x, outs) in enumerate(ops):
            assert 1 <= len(ws) <= 4 and len(ws) == len(outs)
            arr[i].count = len(ws)
            for j, (w, o) in enumerate(zip(ws, outs)):
                assert w.is_cuda and o.is_cuda and o.is_contiguous() and w.stride(-1) == 1
                arr[i].w_dev[j] = w.data_ptr()
                arr[i].M[j] = w.shape[0]
                arr[i].c_dev[j] = o.data_ptr()
            arr[i].x_dev = x.data_ptr()
            arr[i].K = x.shape[-1]
            arr[i].w_stride_bytes = ws[0].stride(0) * ws[0].element_size()
        self.h = C.c_void_p()
        check(L.lammcu_chain_create(w_type, len(ops), arr, C.byref(self.h)), "lammcu_chain_create")

    def run(self):
        check(lib().lammcu_chain_run_dev(self.h, _stream_ptr()), "lammcu_chain_run_dev")

    def timed_out(self):
        v = C.c_int(0)
        check(lib().lammcu_chain_status(self.h, C.byref(v)), "lammcu_chain_status")
        return bool(v.value)

    def close(self):
        if self.h:
            lib().lammcu_chain_destroy(self.h)
            self.h = None


def make_tensor_2d(ttype, ne0, ne1, data_ptr, nb1=None):
    t = GgmlTensor()
    t.type = ttype
    t.ne[0], t.ne[1], t.ne[2], t.ne[3] = ne0, ne1, 1, 1
    ts, bs = type_size(ttype), blck_size(ttype)
    t.nb[0] = ts
    t.nb[1] = nb1 if nb1 is not None else ts * (ne0 // bs)
    t.nb[2] = t.nb[1] * ne1
    t.nb[3] = t.nb[2]
    t.data = data_ptr
    return t


def lamm_mul_mat_host(w_type, w_np, x_np, nth=1, wdata=None, task=TASK_COMPUTE, src0_kind="leaf"):
    """Build ggml_tensor / ggml_compute_params by hand and drive lamm_can_mul_mat +
    lamm_mul_mat exactly as ggml.c:10858-10863 does, once per thread.
    w_np: uint8 [M, row bytes] (or float32 [M, K] for F32); x_np: float32 [N, K].
    src0_kind: "leaf" (a model weight: op NONE, no view_src), "view" (a view of another tensor, e.g. of
    the K cache), "op" (the result of a graph node), "param" (a trainable leaf).
    Returns (can, C[N, M] float32)."""
    import numpy as np
    L = lib()
    m = w_np.shape[0]
    n, k = x_np.shape
    c = np.zeros((n, m), np.float32)
    src0 = make_tensor_2d(w_type, k, m, w_np.ctypes.data, w_np.strides[0])
    src1 = make_tensor_2d(F32, k, n, x_np.ctypes.data, x_np.strides[0])
    dst = make_tensor_2d(F32, m, n, c.ctypes.data)
    dst.src[0] = C.pointer(src0)
    dst.src[1] = C.pointer(src1)
    if src0_kind == "view":
        parent = make_tensor_2d(w_type, k, m, w_np.ctypes.data, w_np.strides[0])
        src0.view_src = C.pointer(parent)
        src0.op = 32                                        # GGML_OP_VIEW (any op != NONE)
    elif src0_kind == "op":
        src0.op = 24                                        # a computed tensor (GGML_OP_MUL_MAT)
    elif src0_kind == "param":
        src0.flags = 4                                      # GGML_TENSOR_FLAG_PARAM
    else:
        assert src0_kind == "leaf"
    can = True
    for ith in range(nth):
        p = GgmlComputeParams(task, ith, nth, wdata.nbytes if wdata is not None else 0,
                              wdata.ctypes.data if wdata is not None else None)
        if not L.lamm_can_mul_mat(C.byref(p), C.byref(dst)):
            can = False
            break
        L.lamm_mul_mat(C.byref(p), C.byref(dst))
    return can, c


I32 = 26                                                    # GGML_TYPE_I32, ggml.h:367


def mul_mat_id_plan(ids_np, id_, n_as):
    """lammcu_mul_mat_id_plan on an int32 [N, n_ids] array -> (counts[n_as], order[N]) or raises LammError.
    Pure host logic: works without a device."""
    import numpy as np
    ids_np = np.ascontiguousarray(ids_np, np.int32)
    n = ids_np.shape[0]
    counts = np.zeros(n_as, np.int64)
    order = np.zeros(max(n, 1), np.int64)
    rc = lib().lammcu_mul_mat_id_plan(ids_np.ctypes.data, ids_np.shape[1], id_, n, n_as, counts.ctypes.data,
                                      order.ctypes.data)
    if rc != 0:
        raise LammError(f"lammcu_mul_mat_id_plan: rc {rc}")
    return counts, order[:n]


def lamm_mul_mat_id_host(w_type, w_list, ids_np, id_, x_np, nth=1):
    """Drive lamm_can_mul_mat_id + lamm_mul_mat_id on hand-built ggml tensors the way a hook at the top of
    ggml_compute_forward_mul_mat_id (ggml.c:10992) would.  w_list: expert matrices (uint8 [M, row bytes], or
    float32 [M, K]); ids_np: int32 [N, n_ids]; x_np: float32 [N, K].  Returns (can, C[N, M])."""
    import numpy as np
    L = lib()
    m = w_list[0].shape[0]
    n, k = x_np.shape
    ids_np = np.ascontiguousarray(ids_np, np.int32)
    c = np.zeros((n, m), np.float32)
    tid = make_tensor_2d(I32, ids_np.shape[1], n, ids_np.ctypes.data, ids_np.strides[0])
    tid.nb[0] = 4
    src1 = make_tensor_2d(F32, k, n, x_np.ctypes.data, x_np.strides[0])
    dst = make_tensor_2d(F32, m, n, c.ctypes.data)
    dst.src[0] = C.pointer(tid)
    dst.src[1] = C.pointer(src1)
    experts = [make_tensor_2d(w_type, k, m, w.ctypes.data, w.strides[0]) for w in w_list]
    for e, t in enumerate(experts):
        dst.src[2 + e] = C.pointer(t)
    dst.op_params[0] = id_
    dst.op_params[1] = len(w_list)
    can = True
    for ith in range(nth):
        p = GgmlComputeParams(TASK_COMPUTE, ith, nth, 0, None)
        if not L.lamm_can_mul_mat_id(C.byref(p), C.byref(dst)):
            can = False
            break
        L.lamm_mul_mat_id(C.byref(p), C.byref(dst))
    return can, c


class Gguf:
    """A memory-mapped GGUF file (include/lammcu.h, lammcu_gguf_*).  Parsing needs no device."""

    def __init__(self, path):
        self.h = C.c_void_p()
        if lib().lammcu_gguf_open(str(path).encode(), C.byref(self.h)) != 0:
            raise LammError(f"lammcu_gguf_open({path}): {lib().lammcu_last_error().decode()}")

    def close(self):
        if self.h:
            lib().lammcu_gguf_close(self.h)
            self.h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __len__(self):
        return int(lib().lammcu_gguf_tensor_count(self.h))

    @property
    def alignment(self):
        return int(lib().lammcu_gguf_alignment(self.h))

    @property
    def data_offset(self):
        return int(lib().lammcu_gguf_data_offset(self.h))

    def tensor(self, i):
        """-> (name, ggml type, n_dims, ne[4], bytes as a numpy uint8 COPY)"""
        import numpy as np
        name, t, nd = C.c_char_p(), C.c_int(), C.c_int()
        ne = (C.c_int64 * 4)()
        data, nb = C.c_void_p(), C.c_uint64()
        if lib().lammcu_gguf_tensor(self.h, i, C.byref(name), C.byref(t), C.byref(nd), ne, C.byref(data), C.byref(nb)) != 0:
            raise LammError("lammcu_gguf_tensor: bad index")
        raw = np.ctypeslib.as_array(C.cast(data, C.POINTER(C.c_uint8)), shape=(nb.value,)).copy() if nb.value else \
            np.zeros(0, np.uint8)
        return name.value.decode(errors="replace"), t.value, nd.value, list(ne), raw

    def find(self, name):
        return int(lib().lammcu_gguf_find(self.h, name.encode()))

    def get_u32(self, key):
        v = C.c_uint32()
        return int(v.value) if lib().lammcu_gguf_get_u32(self.h, key.encode(), C.byref(v)) == 0 else None

    def get_f32(self, key):
        v = C.c_float()
        return float(v.value) if lib().lammcu_gguf_get_f32(self.h, key.encode(), C.byref(v)) == 0 else None

    def get_str(self, key):
        n = lib().lammcu_gguf_get_str(self.h, key.encode(), None, 0)
        if n < 0:
            return None
        buf = C.create_string_buffer(n + 1)
        lib().lammcu_gguf_get_str(self.h, key.encode(), buf, n + 1)
        return buf.value.decode()

    def rows(self, i, row_begin, row_end):
        """Bytes of rows [row_begin, row_end) of tensor i (a rank's shard) as a numpy uint8 COPY."""
        import numpy as np
        data, nb = C.c_void_p(), C.c_uint64()
        if lib().lammcu_gguf_rows(self.h, i, row_begin, row_end, C.byref(data), C.byref(nb)) != 0:
            raise LammError("lammcu_gguf_rows: bad tensor or row range")
        return np.ctypeslib.as_array(C.cast(data, C.POINTER(C.c_uint8)), shape=(nb.value,)).copy()

    def upload_rows(self, i, row_begin, row_end):
        """Register rows [row_begin, row_end) of 2-D tensor i on the device -> weight handle (c_void_p)."""
        w = C.c_void_p()
        if lib().lammcu_gguf_upload_rows(self.h, i, row_begin, row_end, C.byref(w)) != 0:
            raise LammError(f"lammcu_gguf_upload_rows: {lib().lammcu_last_error().decode()}")
        return w
