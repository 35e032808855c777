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
    ap.add_argument("--siblings", type=int, default=0,
                    help="3: three weights called with the same row per token (wq, wk, wv): time line of the speculative jobs")
    a = ap.parse_args()
    if a.siblings:
        return siblings(a)
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


def siblings(a):
    """A group of `siblings` weights called one after the other with the SAME row, a fresh row per group (ggml's
    wq / wk / wv): after the first group the followers' jobs are posted behind the first weight's job.  Per position in
    the group: the host's wall clock per call, and the device-side distance between consecutive jobs' "chunk written"
    stamps (for a follower = everything its job costs on the device: accept in L2, first weights, stream, write-back)."""
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
    ws = [bench.synth_weight(torch, t, a.m, a.k, gen, torch.device("cuda", 0)).cpu().numpy() for _ in range(a.siblings)]
    rng = np.random.default_rng(0)
    groups = 64 // a.siblings
    wall = np.zeros((groups, a.siblings))
    prof = (C.c_double * 8)()
    lib.lammcu_host_profile(prof, 1)
    for g in range(groups):
        x = rng.standard_normal((1, a.k)).astype(np.float32)
        for j, w in enumerate(ws):
            t0 = time.perf_counter()
            can, _c = L.lamm_mul_mat_host(t, w, x, nth=1)
            wall[g, j] = (time.perf_counter() - t0) * 1e6
            assert can
    lib.lammcu_host_profile(prof, 0)
    out = (C.c_ulonglong * 512)()
    lib.lammcu_server_trace.argtypes = [C.POINTER(C.c_ulonglong)]
    assert lib.lammcu_server_trace(out) == 0, L.lib().lammcu_last_error()
    tr = np.array(list(out), dtype=np.float64).reshape(64, 8)
    tr = tr[(tr[:, 0] > 0) & (tr[:, 4] > 0)]
    tr = tr[np.argsort(tr[:, 0])]
    print(f"{a.fmt} {a.m}x{a.k}, groups of {a.siblings} weights on one row: {int(prof[7])} of {int(prof[5])} calls answered by a "
          f"speculative job (python ctypes included in the wall clock)")
    for j in range(a.siblings):
        print(f"  call {j} of the group: host wall median {np.median(wall[4:, j]):6.2f} us")
    follower = (tr[:, 1] - tr[:, 0]) < 1000.0             # no X pull: the job reuses the staged row
    d_prev = (tr[1:, 4] - tr[:-1, 4]) / 1e3               # previous job's chunk written -> this job's
    f = follower[1:]
    print(f"  device, jobs that reuse the row ({int(f.sum())}): previous job's chunk written -> their chunk written median "
          f"{np.median(d_prev[f]):5.2f} us; job seen -> chunk written median {np.median((tr[follower, 4] - tr[follower, 0]) / 1e3):5.2f} us")
    print(f"  device, jobs that carry a row ({int((~follower).sum())}): job seen -> X republished {np.median((tr[~follower, 1] - tr[~follower, 0]) / 1e3):5.2f} us, "
          f"-> quantized {np.median((tr[~follower, 2] - tr[~follower, 0]) / 1e3):5.2f} us, -> chunk written {np.median((tr[~follower, 4] - tr[~follower, 0]) / 1e3):5.2f} us")

if __name__ == "__main__":
    main()
