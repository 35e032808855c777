"""Device time line of resident-server jobs (diagnostics, run on the GPU box):

    LAMM_SERVER_TRACE=1 python tools/trace_server.py [--m 4096 --k 4096]

Posts 64 N == 1 jobs through lamm_mul_mat on host tensors and prints, from the %globaltimer stamps the
server kernel leaves per job (lammcu_server_trace): job seen -> X republished -> X quantized -> CTA 0's
rows done -> CTA 0's chunk written, plus the host's wall clock per call.  Medians over the last 48 jobs."""
import argparse
import ctypes as C
import os
import sys
import time

os.environ.setdefault("LAMM_SERVER_TRACE", "1")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, default=4096)
    ap.add_argument("--k", type=int, default=4096)
    ap.add_argument("--fmt", default="q4_1")
    a = ap.parse_args()
    import torch
    import lamm_b200 as L
    import bench
    torch.cuda.set_device(0)
    L.init(0)
    lib = L.lib()
    assert lib.lammcu_set_host_mode(1) == 0
    t = bench.TYPE_IDS[a.fmt]
    gen = torch.Generator(device="cuda")
    gen.manual_seed(1)
    w = bench.synth_weight(torch, t, a.m, a.k, gen, torch.device("cuda", 0)).cpu().numpy()
    rng = np.random.default_rng(0)
    wall = []
    for _ in range(64):
        x = rng.standard_normal((1, a.k)).astype(np.float32)
        t0 = time.perf_counter()
        can, _c = L.lamm_mul_mat_host(t, w, x, nth=1)
        wall.append((time.perf_counter() - t0) * 1e6)
        assert can
    out = (C.c_ulonglong * 512)()
    lib.lammcu_server_trace.argtypes = [C.POINTER(C.c_ulonglong)]
    assert lib.lammcu_server_trace(out) == 0, L.lib().lammcu_last_error()
    tr = np.array(list(out), dtype=np.float64).reshape(64, 8)
    tr = tr[(tr[:, 0] > 0) & (tr[:, 4] > 0)]
    tr = tr[np.argsort(tr[:, 0])]                      # rows are indexed by job & 63: make them chronological
    d = tr[-48:, 1:5] - tr[-48:, 0:1]
    names = ("X republished", "X quantized", "CTA0 rows done", "CTA0 chunk written")
    print(f"{a.fmt} {a.m}x{a.k}: host wall/call median {np.median(wall[16:]):.2f} us "
          f"(python ctypes included)")
    for i, n in enumerate(names):
        print(f"  job seen -> {n:20s} median {np.median(d[:, i]) / 1e3:6.2f} us   p90 {np.percentile(d[:, i], 90) / 1e3:6.2f} us")
    gap = np.diff(tr[-48:, 0])
    print(f"  job seen -> next job seen       median {np.median(gap) / 1e3:6.2f} us")


if __name__ == "__main__":
    main()
