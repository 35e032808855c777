"""One prefill GEMM shape held for seconds: burst rate vs the rate under the board's power cap (diagnostics):

    python tools/gemm_sustained.py q4_1 13824 5120 2048 [seconds]

The first 5 launches after a second of idle run at the maximum SM clock; a tensor-bound kernel repeated for seconds does
not (sw_power_cap).  Prints both rates and nvidia-smi's clocks / power over the sustained leg (bench.ClockSampler)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import lamm_b200 as L
    import bench
    fmt, m, k, n = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    seconds = float(sys.argv[5]) if len(sys.argv) > 5 else 2.0
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    L.init(0)
    t = bench.TYPE_IDS[fmt]
    gen = torch.Generator(device=dev)
    gen.manual_seed(1)
    w = bench.synth_weight(torch, t, m, k, gen, dev)
    x = torch.randn((n, k), generator=gen, device=dev)
    out = torch.empty((n, m), device=dev)
    ops = 2.0 * m * n * k

    def timed(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            L.mul_mat(t, w, x, out=out)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / reps
    timed(3)
    time.sleep(1.0)
    burst = timed(5)
    reps = max(10, int(seconds * 1e6 / burst))
    timed(reps // 2)                                           # settle clocks and power
    smp = bench.ClockSampler(0)
    smp.start()
    sus = timed(reps)
    clocks = smp.stop()
    print(f"{fmt} {m}x{k} N={n}: burst {burst:.1f} us = {ops / burst / 1e6:.1f} TOPS; sustained {sus:.1f} us = "
          f"{ops / sus / 1e6:.1f} TOPS over {reps} launches; clocks {clocks}")


if __name__ == "__main__":
    main()
