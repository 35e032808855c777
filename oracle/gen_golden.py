"""Generate tests/golden/*.npz from the REAL reference (oracle/_ref/libref.so).

Run in the build container only (needs /root/reference):  python oracle/gen_golden.py
The reference ships no golden vectors of its own, so these fixtures are
outputs of the reference itself on seeded inputs:
  wq        weights quantized by ggml_quantize_chunk (ggml-quants.c *_reference)
  x         f32 activations: Gaussian, plus an all-zero row, a row with zero blocks,
            and a row with +-amax ties / exact .5 rounding cases
  xq        quantize_row_q8_{0,1,K}_reference(x)   (0xAB pre-fill shows unwritten bytes)
  c_lamm    lamm_mul_mat called directly on (wq, xq)      [wrong for Q8_0 on AVX2, 9-2]
  c_vecdot  ggml_vec_dot_* on (wq, xq)
  c_graph   the whole path through ggml_graph_compute (AVX2 INIT quantizer)
Shapes follow test/test_correctness.py's awkward M=33, N=18 (src/la-benchmark-matmult.cpp:176-178).
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from oracle_lib import (F16, F32, KQ_TYPES, Q2_K, Q8_0, Q8_1, Q8_K, TYPE_NAMES, VEC_DOT, W_TYPES, Ref, build_ref)

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def make_x(rng, n, k):
    x = rng.standard_normal((n, k)).astype(np.float32)
    x[1] = 0.0                                   # all-zero row: d = 0, id = 0
    x[2, : k // 2] = 0.0                         # zero blocks inside a live row
    x[3, 0::7] = 2.5                             # many equal maxima, first one must win
    x[3, 5] = -2.5
    x[4] = np.round(x[4] * 4) / 4                # values landing on .5 after scaling
    x[4, 0] = 127.0 / 2                          # d = 0.5 -> x*id hits exact halves
    x[5] *= 1e-6                                 # fp16-subnormal d
    x[6] *= 3e3                                  # large d
    return x


def main():
    assert build_ref(), "needs /root/reference"
    r = Ref()
    os.makedirs(OUT, exist_ok=True)
    rng = np.random.default_rng(20240314)
    for (m, n, k, types) in [(33, 18, 512, W_TYPES), (33, 18, 4096, [3])]:
        w = rng.standard_normal((m, k)).astype(np.float32)
        x = make_x(rng, n, k)
        for t in types:
            name = TYPE_NAMES[t]
            if t == F32:
                wq = w.copy().view(np.uint8).reshape(m, k * 4)
                c_graph = r.graph_mul_mat(t, wq, x)
                np.savez_compressed(os.path.join(OUT, f"mulmat_{name}_m{m}_n{n}_k{k}.npz"),
                                    w_type=t, wq=wq, x=x, c_graph=c_graph)
                continue
            vt = VEC_DOT[t]
            wq = r.quantize_weights(t, w)
            xq = r.quantize_act(vt, x, prefill=0xAB)
            xq0 = r.quantize_act(vt, x, prefill=0)
            c_lamm = r.lamm_mul_mat(t, wq, x, xq0)
            c_vecdot = r.vec_dot(t, wq, xq0, k)
            c_graph = r.graph_mul_mat(t, wq, x)
            np.savez_compressed(os.path.join(OUT, f"mulmat_{name}_m{m}_n{n}_k{k}.npz"),
                                w_type=t, wq=wq, x=x, xq=xq, c_lamm=c_lamm, c_vecdot=c_vecdot,
                                c_graph=c_graph, w_deq=r.dequantize(t, wq, k))
            print(name, m, n, k, "lamm-vecdot", float(np.abs(c_lamm - c_vecdot).max()),
                  "graph-vecdot", float(np.abs(c_graph - c_vecdot).max()))
    # activation quantizer fixtures on their own (bigger, more edge rows)
    x = make_x(rng, 12, 2048)
    x[7, 256:512] = 0.0
    x[8] = np.float32(1.0)                        # constant block
    x[9, ::2] *= -1
    np.savez_compressed(os.path.join(OUT, "quantize_act_n12_k2048.npz"), x=x,
                        q8_0=r.quantize_act(Q8_0, x, prefill=0xAB),
                        q8_1=r.quantize_act(Q8_1, x, prefill=0xAB),
                        q8_k=r.quantize_act(Q8_K, x, prefill=0xAB))
    # SURVEY 8f-1 K-quants (own generator so the fixtures above stay byte-identical): lamm's gate refuses
    # these types, so the reference behaviour is ggml's own path (c_graph) / ggml_vec_dot_q*_K_q8_K
    rng = np.random.default_rng(20240315)
    m, n, k = 33, 18, 1024
    w = rng.standard_normal((m, k)).astype(np.float32)
    x = make_x(rng, n, k)
    for t in KQ_TYPES:
        name = TYPE_NAMES[t]
        wq = r.quantize_weights(t, w)
        xq = r.quantize_act(Q8_K, x, prefill=0xAB)
        xq0 = r.quantize_act(Q8_K, x, prefill=0)
        c_vecdot = r.vec_dot(t, wq, xq0, k)
        c_graph = r.graph_mul_mat(t, wq, x)
        np.savez_compressed(os.path.join(OUT, f"mulmat_{name}_m{m}_n{n}_k{k}.npz"),
                            w_type=t, wq=wq, x=x, xq=xq, c_vecdot=c_vecdot, c_graph=c_graph,
                            w_deq=r.dequantize(t, wq, k))
        print(name, m, n, k, "graph-vecdot", float(np.abs(c_graph - c_vecdot).max()))
    # SURVEY 8f-2, oracle step: F16 weights through stock ggml (lamm's gate refuses F16).  Own generator and
    # an own file-name prefix: the CUDA path has no F16 kernel yet, and the GPU tests sweep mulmat_*.npz
    rng = np.random.default_rng(20240316)
    m, n, k = 33, 18, 4096
    w = rng.standard_normal((m, k)).astype(np.float32)
    x = make_x(rng, n, k)
    x[6] *= np.float32(1e-2)                      # keep every row inside the fp16 range (make_x scales it by 3e3)
    wq = r.quantize_weights(F16, w)
    np.savez_compressed(os.path.join(OUT, f"next_f16_m{m}_n{n}_k{k}.npz"), w_type=F16, wq=wq, x=x,
                        c_graph=r.graph_mul_mat(F16, wq, x))
    # llm_build_norm and llm_build_ffn's silu(gate) * up in front of a Q4_1 matmul, through stock ggml's graph
    # (one thread: SURVEY 9-3).  y = the f32 activation the matmul consumed; Gaussian rows only (no quantizer ties: the
    # x86 INIT quantizer rounds halves to even)
    rng = np.random.default_rng(20240317)
    m, n, k = 33, 6, 2048
    wq = r.quantize_weights(3, rng.standard_normal((m, k)).astype(np.float32))
    x = (rng.standard_normal((n, k)) * 3).astype(np.float32)
    x[1] = 0.0
    norm_w = (1.0 + 0.1 * rng.standard_normal(k)).astype(np.float32)
    c1, y1 = r.graph_xop_mul_mat(1, 3, wq, x, norm_w, 1e-5)
    gate = (rng.standard_normal((n, k)) * 4).astype(np.float32)
    up = rng.standard_normal((n, k)).astype(np.float32)
    c2, y2 = r.graph_xop_mul_mat(2, 3, wq, gate, up)
    np.savez_compressed(os.path.join(OUT, f"xop_q4_1_m{m}_n{n}_k{k}.npz"), w_type=3, wq=wq, x=x, norm_w=norm_w,
                        eps=np.float32(1e-5), y_norm=y1, c_norm=c1, gate=gate, up=up, y_silu=y2, c_silu=c2)
    # a GGUF file written by the reference's own writer (key/values of every kind + 9 tensors)
    from oracle_lib import Oracle
    from test_gguf import FIXTURE, fixture_tensors
    r.gguf_write(FIXTURE, fixture_tensors(Oracle()), alignment=0)
    print("gguf fixture", os.path.getsize(FIXTURE), "bytes")


if __name__ == "__main__":
    main()
