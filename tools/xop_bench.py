"""what fusing the op in front of the matmul saves, 7B Q4_1 shapes, CUDA events, 200 iterations each
(weights of one shape: L2-warm, so the matmul part is the same small constant on both sides).

  unfused = the op as its own launch writing f32 (our own kernel, VT = F32), then lammcu_mul_mat[_batch]_dev
  fused   = lammcu_mul_mat[_batch]_xop_dev

    python tools/xop_bench.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(torch, fn, iters=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def main():
    import torch
    import lamm_b200 as L
    import bench
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    L.init(0)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1)
    t = 3
    k, f = 4096, 11008
    wq, wk, wv = (bench.synth_weight(torch, t, 4096, k, gen, dev) for _ in range(3))
    wg, wu = (bench.synth_weight(torch, t, f, k, gen, dev) for _ in range(2))
    wd = bench.synth_weight(torch, t, k, f, gen, dev)
    nw = torch.ones((k,), device=dev)
    for n in (1, 512):
        x = torch.randn((n, k), generator=gen, device=dev)
        gate = torch.randn((n, f), generator=gen, device=dev)
        up = torch.randn((n, f), generator=gen, device=dev)
        norm, silu = L.rms_norm_op(nw, 1e-5), L.silu_mul_op(up)
        if n == 1:
            outs = [torch.empty((1, 4096), device=dev) for _ in range(3)]
            us_a = timed(torch, lambda: L.mul_mat_batch(t, [wq, wk, wv], L.quantize_act_xop(norm, x, 0).view(torch.float32), outs))
            us_b = timed(torch, lambda: L.mul_mat_batch_xop(norm, t, [wq, wk, wv], x, outs))
            print(f"N=1 attn_norm + q/k/v      : unfused {us_a:7.2f} us   fused {us_b:7.2f} us")
            outs = [torch.empty((1, f), device=dev) for _ in range(2)]
            us_a = timed(torch, lambda: L.mul_mat_batch(t, [wg, wu], L.quantize_act_xop(norm, x, 0).view(torch.float32), outs))
            us_b = timed(torch, lambda: L.mul_mat_batch_xop(norm, t, [wg, wu], x, outs))
            print(f"N=1 ffn_norm + gate/up     : unfused {us_a:7.2f} us   fused {us_b:7.2f} us")
        else:
            out = torch.empty((n, 4096), device=dev)
            us_a = timed(torch, lambda: L.mul_mat(t, wq, L.quantize_act_xop(norm, x, 0).view(torch.float32), out=out), 50)
            us_b = timed(torch, lambda: L.mul_mat_xop(norm, t, wq, x, out=out), 50)
            print(f"N={n} attn_norm + wq       : unfused {us_a:7.2f} us   fused {us_b:7.2f} us")
        out = torch.empty((n, k), device=dev)
        it = 200 if n == 1 else 50
        us_a = timed(torch, lambda: L.mul_mat(t, wd, L.quantize_act_xop(silu, gate, 0).view(torch.float32), out=out), it)
        us_b = timed(torch, lambda: L.mul_mat_xop(silu, t, wd, gate, out=out), it)
        print(f"N={n} silu(gate)*up + down  : unfused {us_a:7.2f} us   fused {us_b:7.2f} us")


if __name__ == "__main__":
    main()
