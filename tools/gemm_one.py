"""One prefill GEMM shape, a few launches (for ncu):  python tools/gemm_one.py q2_k 4096 4096 512 [exact-mode]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import lamm_b200 as L
    import bench
    fmt, m, k, n = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    mode = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    L.init(0)
    L.lib().lammcu_set_gemm_exact(mode)
    t = bench.TYPE_IDS[fmt]
    gen = torch.Generator(device=dev)
    gen.manual_seed(1)
    w = bench.synth_weight(torch, t, m, k, gen, dev)
    x = torch.randn((n, k), generator=gen, device=dev)
    out = torch.empty((n, m), device=dev)
    for _ in range(3):
        L.mul_mat(t, w, x, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        L.mul_mat(t, w, x, out=out)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 5
    print(f"{fmt} {m}x{k} N={n} mode {mode}: {us:.1f} us, {2.0 * m * n * k / us / 1e6:.1f} TOPS (L2-warm), "
          f"finite={bool(torch.isfinite(out).all())}")


if __name__ == "__main__":
    main()
