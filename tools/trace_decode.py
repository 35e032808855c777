"""In-kernel time line of the single-launch decode kernel inside a chain (diagnostics, run on the GPU box):

    python tools/trace_decode.py [--fmt q4_1] [--m 4096 --k 4096] [--chain 12]

Launches `chain` data-dependent mul_mats (each one's output row is the next one's activation row, PDL between
them, distinct weights) with lammcu_set_trace armed and prints, per launch, medians over the CTAs of the
stamps decode16_kernel leaves (lamm_decode.cu: trace_stamp): [0] %globaltimer at CTA start; [1..7] clock64 at
start / after griddepcontrol.wait / activations quantized / CTA done / lane setup done (about to wait for the
weights) / first weights landed / first row's dot done.  clock64 deltas are converted with the SM clock."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--fmt", default="q4_1")
    ap.add_argument("--m", type=int, default=4096)
    ap.add_argument("--k", type=int, default=4096)
    ap.add_argument("--chain", type=int, default=12)
    ap.add_argument("--mhz", type=float, default=1965.0)
    a = ap.parse_args()
    assert a.m == a.k, "a chain needs square matrices (output row = next activation row)"
    import torch
    import lamm_b200 as L
    import bench
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    L.init(0)
    lib = L.lib()
    t = bench.TYPE_IDS[a.fmt]
    gen = torch.Generator(device=dev)
    gen.manual_seed(1)
    ws = [bench.synth_weight(torch, t, a.m, a.k, gen, dev) for _ in range(a.chain)]
    bufs = [torch.randn((1, a.k), generator=gen, device=dev) for _ in range(a.chain + 1)]
    sms = lib.lammcu_sm_count()
    trace = torch.zeros((a.chain * sms * 8,), dtype=torch.int64, device=dev)

    def run():
        for i, w in enumerate(ws):
            L.mul_mat(t, w, bufs[i], out=bufs[i + 1])
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        run()
    torch.cuda.synchronize()
    # the trace pointers are baked into the launches at capture time: replaying the graph (back-to-back
    # launches, PDL overlap as in bench.py) overwrites the same slots, the last replay is what is read
    lib.lammcu_set_trace(trace.data_ptr(), trace.numel())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        run()
    lib.lammcu_set_trace(None, 0)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    tr = trace.cpu().numpy().reshape(a.chain, sms, 8).astype(np.float64)
    cyc = 1e3 / a.mhz                                  # ns per cycle
    print(f"{a.fmt} {a.m}x{a.k}, {a.chain} chained launches, {sms} CTAs each; ns, medians over CTAs")
    print("launch  start-prev_start  wait_returns  x_quantized  lane_setup  first_weights  row0  cta_done")
    prev = None
    for i in range(a.chain):
        r = tr[i]
        live = r[:, 4] > 0
        r = r[live]
        start = np.median(r[:, 0])
        rel = lambda e: np.median((r[:, e] - r[:, 1]) * cyc)
        print(f"{i:5d} {(start - prev) if prev else 0:16.0f} {rel(2):13.0f} {rel(3):12.0f} {rel(5):14.0f} {rel(6):5.0f} {rel(7):5.0f} {rel(4):9.0f}")
        prev = start


if __name__ == "__main__":
    main()
