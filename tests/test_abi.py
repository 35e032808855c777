"""The C-ABI boundary on a CPU-only machine: the library loads, exports every symbol the
headers declare, the type table matches ggml's, the struct layouts match ggml.h, and compute
entry points fail loudly (no CPU fallback) when there is no device."""
import ctypes as C
import json
import os
import re
import subprocess
import tempfile

import numpy as np
import pytest
from oracle_lib import BLCK, TSIZE, VEC_DOT

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return set(re.findall(r"\b(lammcu_[a-z0-9_]+|lamm_[a-z_]+)\s*\(", txt))


def test_exports_every_declared_symbol(lamm):
    L = lamm.lib()
    declared = _declared("lammcu.h") | _declared("lamm_b200.h")
    assert {"lamm_can_mul_mat", "lamm_mul_mat", "lamm_get_opt_level"} <= declared
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/ but not exported"
    assert declared == set(lamm.EXPORTS)


def test_type_table(lamm):
    for t, bs in BLCK.items():
        assert lamm.blck_size(t) == bs and lamm.type_size(t) == TSIZE[t]
        assert lamm.row_size(t, 4096) == 4096 // bs * TSIZE[t]
    for t, vt in VEC_DOT.items():
        assert lamm.vec_dot_type(t) == vt and lamm.lib().lammcu_type_supported(t) == 1
    for t in (9, 15, 16):                         # Q8_1, Q8_K, IQ2_XXS: not weight types of the path
        assert lamm.lib().lammcu_type_supported(t) == 0


def test_opt_level_default(lamm):
    assert lamm.lib().lamm_get_opt_level() == int(os.environ.get("LAMM_OPT_LEVEL", 3))


OFFSET_PROG = r"""
#include <stdio.h>
#include <stddef.h>
%s
#define F(s, f) printf("\"" #s "." #f "\": [%%zu, %%zu],\n", offsetof(struct s, f), sizeof(((struct s*)0)->f))
int main(void) {
  printf("{\n");
  F(ggml_tensor, type); F(ggml_tensor, backend); F(ggml_tensor, buffer); F(ggml_tensor, ne);
  F(ggml_tensor, nb); F(ggml_tensor, op); F(ggml_tensor, op_params); F(ggml_tensor, flags);
  F(ggml_tensor, grad); F(ggml_tensor, src); F(ggml_tensor, view_src); F(ggml_tensor, view_offs);
  F(ggml_tensor, data); F(ggml_tensor, name); F(ggml_tensor, extra);
  F(ggml_compute_params, type); F(ggml_compute_params, ith); F(ggml_compute_params, nth);
  F(ggml_compute_params, wsize); F(ggml_compute_params, wdata);
  printf("\"sizeof.ggml_tensor\": [%%zu, 0], \"sizeof.ggml_compute_params\": [%%zu, 0],\n",
         sizeof(struct ggml_tensor), sizeof(struct ggml_compute_params));
  printf("\"enum\": [%%d, %%d]\n}\n", (int)GGML_TYPE_Q8_K, (int)GGML_TASK_TYPE_COMPUTE);
  return 0;
}
"""


def _layout(includes, cflags):
    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "o.c")
        open(src, "w").write(OFFSET_PROG % includes)
        exe = os.path.join(d, "o")
        subprocess.check_call(["gcc", "-std=c11", "-w"] + cflags + [src, "-o", exe])
        return json.loads(subprocess.check_output([exe]).decode())


def test_struct_layout_matches_ggml(lamm):
    """include/lamm_b200.h's restated structs == llama.cpp-b2430/ggml.h:552-590,668-677.
    Against the real header where /root/reference exists, else the committed snapshot."""
    ours = _layout('#include "lamm_b200.h"', ["-I", os.path.join(ROOT, "include")])
    snap = os.path.join(ROOT, "tests", "golden", "ggml_abi_b2430.json")
    ref_dir = "/root/reference/llama.cpp-b2430"
    if os.path.isdir(ref_dir):
        theirs = _layout('#include "ggml.h"', ["-I", ref_dir])
        if not os.path.exists(snap):
            json.dump(theirs, open(snap, "w"), indent=1, sort_keys=True)
        assert theirs == json.load(open(snap))
    else:
        theirs = json.load(open(snap))
    assert ours == theirs
    # and the ctypes mirror used by the Python binding
    assert C.sizeof(lamm.GgmlTensor) == theirs["sizeof.ggml_tensor"][0]
    assert C.sizeof(lamm.GgmlComputeParams) == theirs["sizeof.ggml_compute_params"][0]
    for f in ("ne", "nb", "src", "data"):
        assert getattr(lamm.GgmlTensor, f).offset == theirs[f"ggml_tensor.{f}"][0]


def _no_device(lamm):
    return lamm.lib().lammcu_device_count() <= 0


def test_gate_semantics_without_running(lamm):
    """lamm_can_mul_mat: false outside COMPUTE, for unsupported types, bad K (src/loongarch_matmul.cpp:10-62)."""
    w = np.zeros((4, 20 * 8), np.uint8)
    x = np.zeros((2, 256), np.float32)
    for task in (lamm.TASK_INIT, lamm.TASK_FINALIZE):
        can, _ = lamm.lamm_mul_mat_host(lamm.Q4_1, w, x, task=task)
        assert can is False
    can, _ = lamm.lamm_mul_mat_host(12, np.zeros((4, 144), np.uint8), x)    # Q4_K: not supported
    assert can is False
    if _no_device(lamm):
        can, _ = lamm.lamm_mul_mat_host(lamm.Q4_1, w, x)
        assert can is False, "without a GPU the gate must send ggml down its own path"


def test_no_cpu_fallback(lamm):
    """Product path must fail loudly when there is no CUDA device."""
    if not _no_device(lamm):
        pytest.skip("a CUDA device is present")
    L = lamm.lib()
    assert L.lammcu_init(0) == -1                                    # LAMMCU_ERR_NO_DEVICE
    buf = np.zeros(4096, np.float32)
    rc = L.lammcu_mul_mat_dev(lamm.Q4_1, buf.ctypes.data, 20, buf.ctypes.data, 32, buf.ctypes.data, 1, 1, 1, 32, None, None)
    assert rc == -1 and b"CUDA" in L.lammcu_last_error() or b"device" in L.lammcu_last_error()
    h = C.c_void_p()
    assert L.lammcu_weight_register(buf.ctypes.data, lamm.Q4_1, 1, 32, 20, 0, 1, C.byref(h)) == -1


def test_every_environment_variable_is_documented():
    """INTEGRATION.md section 6 lists every LAMM_* variable the library reads (and nothing it does not)."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    read = set()
    csrc = os.path.join(root, "la-llama.cpp_b200", "csrc")
    for f in os.listdir(csrc):
        with open(os.path.join(csrc, f)) as fh:
            read |= set(re.findall(r'getenv\("(LAMM_[A-Z0-9_]+)"\)', fh.read()))
    with open(os.path.join(root, "la-llama.cpp_b200", "__init__.py")) as fh:      # the ctypes binding reads LAMM_LIB
        read |= set(re.findall(r'os\.environ\.get\("(LAMM_[A-Z0-9_]+)"\)', fh.read()))
    with open(os.path.join(root, "INTEGRATION.md")) as fh:
        doc = fh.read()
    table = set(re.findall(r"^\| `(LAMM_[A-Z0-9_]+)` \|", doc, flags=re.M))
    assert read, "no getenv found: the scan is broken"
    assert read == table, f"undocumented: {sorted(read - table)}; stale: {sorted(table - read)}"


def test_mul_mat_id_plan_groups_tokens_like_ggml(lamm):
    """lammcu_mul_mat_id_plan (pure host logic) against the row lists ggml_compute_forward_mul_mat_id builds
    in its INIT phase (ggml.c:11066-11078): per-expert counts, tokens in ascending order within an expert."""
    rng = np.random.default_rng(5)
    for n, n_as, n_ids in [(1, 4, 2), (11, 4, 2), (64, 8, 3), (7, 1, 1), (0, 3, 1)]:
        ids = rng.integers(0, n_as, (n, n_ids)).astype(np.int32)
        for id_ in range(n_ids):
            counts, order = lamm.mul_mat_id_plan(ids, id_, n_as)
            want_rows = [[j for j in range(n) if ids[j, id_] == e] for e in range(n_as)]
            assert counts.tolist() == [len(r) for r in want_rows]
            assert order.tolist() == [j for r in want_rows for j in r]
    ids = np.array([[0, 1], [3, 1], [1, 4]], np.int32)
    assert lamm.mul_mat_id_plan(ids, 0, 4)[0].tolist() == [1, 1, 0, 1]
    for bad_id, n_as in [(1, 4), (2, 8), (-1, 8), (0, 3)]:          # id 4 >= n_as, column outside ids, expert 3 >= 3
        with pytest.raises(lamm.LammError):
            lamm.mul_mat_id_plan(ids, bad_id, n_as)
    ids[1, 0] = -1
    with pytest.raises(lamm.LammError):
        lamm.mul_mat_id_plan(ids, 0, 4)


def test_mul_mat_id_gate_without_device(lamm):
    """lamm_can_mul_mat_id: false without a CUDA device (ggml then keeps its own path), for a well-formed node."""
    if not _no_device(lamm):
        pytest.skip("a CUDA device is present")
    w = [np.zeros((4, 20 * 8), np.uint8) for _ in range(3)]          # Q4_1, K = 256
    ids = np.array([[0], [2]], np.int32)
    can, c = lamm.lamm_mul_mat_id_host(lamm.Q4_1, w, ids, 0, np.zeros((2, 256), np.float32))
    assert can is False and not c.any()
