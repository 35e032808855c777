"""In-kernel time line of the persistent decode chain (diagnostics, run on the GPU box):

    python tools/trace_chain.py [--fmt q4_1]

Runs the LLaMA-7B chain of bench.py once with lammcu_set_trace armed and prints, per kind of stage, the medians of
CTA 0's stamps: barrier wait, activation quantize, rows (stream + dot), and the stage period."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--fmt", default="q4_1")
    a = ap.parse_args()
    import torch
    import lamm_b200 as L
    import bench
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    L.init(0)
    lib = L.lib()
    t = bench.TYPE_IDS[a.fmt]
    H, F, V, LAYERS = bench.H, bench.F, bench.V, bench.LAYERS
    gen = torch.Generator(device=dev)
    gen.manual_seed(1)
    mk = lambda m: torch.zeros((1, m), device=dev)
    x0 = torch.randn((1, H), generator=gen, device=dev)
    ops, kinds = [], []
    xin = x0
    for _ in range(LAYERS):
        q, k, v, o, g, u, xo = mk(H), mk(H), mk(H), mk(H), mk(F), mk(F), mk(H)
        w = lambda m, kk: bench.synth_weight(torch, t, m, kk, gen, dev)
        for name, ws, x, outs in (("qkv", [w(H, H), w(H, H), w(H, H)], xin, [q, k, v]), ("wo", [w(H, H)], v, [o]),
                                  ("gate_up", [w(F, H), w(F, H)], o, [g, u]), ("down", [w(H, F)], u, [xo])):
            ops.append((ws, x, outs))
            kinds.append(name)
        xin = xo
    ops.append(([bench.synth_weight(torch, t, V, H, gen, dev)], xin, [mk(V)]))
    kinds.append("head")
    ch = L.Chain(t, ops)
    for _ in range(3):
        ch.run()
    torch.cuda.synchronize()
    sms = lib.lammcu_sm_count()
    trace = torch.zeros((sms * 8,), dtype=torch.int64, device=dev)
    lib.lammcu_set_trace(trace.data_ptr(), trace.numel())
    ch.run()
    lib.lammcu_set_trace(None, 0)
    torch.cuda.synchronize()
    tr = trace.cpu().numpy()[: len(ops) * 4].reshape(len(ops), 4).astype(np.float64)
    total = tr[-1, 3] - tr[0, 0]
    print(f"{a.fmt}: {len(ops)} stages, token {total / 1e3:.1f} us (CTA 0)")
    print("stage      MB   barrier_wait  quantize   rows   period   GB/s(period)")
    for name in ("qkv", "wo", "gate_up", "down", "head"):
        idx = [i for i, kname in enumerate(kinds) if kname == name and 0 < i < len(ops) - 1 or (kname == name == "head" and i > 0)]
        if not idx:
            continue
        mb = sum(w_.shape[0] * w_.shape[1] for w_ in ops[idx[0]][0]) / 1e6
        bw = np.median([tr[i, 1] - tr[i, 0] for i in idx])
        qz = np.median([tr[i, 2] - tr[i, 1] for i in idx])
        rows = np.median([tr[i, 3] - tr[i, 2] for i in idx])
        per = np.median([tr[i, 3] - tr[i - 1, 3] for i in idx])
        print(f"{name:8s} {mb:6.1f} {bw:12.0f} {qz:9.0f} {rows:7.0f} {per:8.0f} {mb * 1e3 / per:10.0f}")
    ch.close()


if __name__ == "__main__":
    main()
