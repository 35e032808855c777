"""lammcu_mul_mat_noalloc_dev (include/lammcu.h; no counterpart in the reference):
the 1 <= N <= 8 quantize + GEMV path with its quantized-activation scratch in a static device array
instead of a cudaMalloc'ed workspace.  Called through the C-ABI and compared with the existing entry point
lammcu_mul_mat_dev (and the oracle) within the tier tolerance, on two calls in a row that share the array.
Collected last (test_zz_*)."""
import ctypes as C

import numpy as np
import pytest
from conftest import make_x, tol_c
from oracle_lib import F32, Q2_K, Q4_1, Q5_0, Q8_0, TYPE_NAMES

pytestmark = pytest.mark.gpu


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _noalloc(gpu, t, wd, xd, m, n, k):
    import torch
    lib = gpu.lib()
    out = torch.full((n, m), float("nan"), dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.lammcu_mul_mat_noalloc_dev(t, _ptr(wd), wd.stride(0) * wd.element_size(), C.cast(_ptr(xd), C.POINTER(C.c_float)),
                                        xd.stride(0), C.cast(_ptr(out), C.POINTER(C.c_float)), m, m, n, k, C.c_void_p(st))
    assert rc == 0, lib.lammcu_last_error()
    return out


@pytest.mark.parametrize("t", [Q4_1, Q5_0, Q8_0, Q2_K])
def test_noalloc_matches_existing_entry_point_twice(gpu, oracle, t):
    import torch
    from test_gpu_parity import dev, weights
    rng = np.random.default_rng(40 + t)
    # two calls in a row through the shared scratch: different shapes, so the second call sees the first one's
    # leftovers in the array (a longer row, then more and shorter rows)
    for (m, n, k) in [(517, 3, 4096), (300, 8, 1024)]:
        w = weights(oracle, t, m, k, seed=t + m)
        x = make_x(rng, 8, k)[:n]
        wd, xd = dev(w), dev(x)
        got = _noalloc(gpu, t, wd, xd, m, n, k)
        ref = gpu.mul_mat(t, wd, xd)                         # the existing entry point (lammcu_mul_mat_dev)
        torch.cuda.synchronize()
        got, ref = got.cpu().numpy(), ref.cpu().numpy()
        want = oracle.mul_mat(t, w, x)
        assert np.isfinite(got).all()
        assert np.abs(got - ref).max() <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{n}x{k} vs lammcu_mul_mat_dev"
        assert np.abs(got - want).max() <= tol_c(want, k), f"{TYPE_NAMES[t]} {m}x{n}x{k} vs oracle"


def test_noalloc_rejects_what_it_cannot_hold(gpu, oracle):
    import torch
    from test_gpu_parity import dev, weights
    lib = gpu.lib()
    k = 1024
    w = dev(weights(oracle, Q4_1, 16, k, seed=1))
    x = dev(np.zeros((9, k), np.float32))
    out = torch.zeros((9, 16), dtype=torch.float32, device="cuda")
    fp = C.POINTER(C.c_float)

    def call(t, n, kk, wt=w):
        return lib.lammcu_mul_mat_noalloc_dev(t, _ptr(wt), kk // 32 * 20, C.cast(_ptr(x), fp), kk, C.cast(_ptr(out), fp),
                                              16, 16, n, kk, None)
    assert call(Q4_1, 9, k) == -4 and b"N <= 8" in lib.lammcu_last_error()          # LAMMCU_ERR_SHAPE
    assert call(Q4_1, 8, 1 << 20) == -4 and b"static array holds" in lib.lammcu_last_error()   # 8 x 1.18 MB > 512 KiB
    assert call(F32, 1, k) == -3                                                     # LAMMCU_ERR_TYPE
    assert call(Q4_1, 8, k) == 0
    torch.cuda.synchronize()
