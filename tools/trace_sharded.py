"""In-kernel time line of the ROW-SHARDED decode kernel inside a chain (diagnostics, run under torchrun on a GPU box):

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/trace_sharded.py [--m 4096] [--chain 12]

Every rank launches `chain` data-dependent sharded mul_mats (M = K; each launch's LL output vector is the next one's
activation row, so every launch waits for the peer stores of all ranks) with lammcu_set_trace armed; rank 0 prints, per
launch, medians over its CTAs of the stamps of decode16_kernel (see tools/trace_decode.py): start, tags of the
activation row seen ("wait_returns" = end of the poll), activations quantized, lane setup, first weights, first row, done."""
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
    ap.add_argument("--chain", type=int, default=12)
    ap.add_argument("--mhz", type=float, default=1965.0)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    import lamm_b200 as L
    import bench
    from lamm_b200 import sharding
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    L.init(local)
    lib = L.lib()
    comm = sharding.Comm(L, dist)
    t = bench.TYPE_IDS[a.fmt]
    gen = torch.Generator(device=dev)
    gen.manual_seed(1)
    lo, hi = sharding.row_shard(a.m, rank, world)
    ws = [bench.synth_weight(torch, t, hi - lo, a.m, gen, dev) for _ in range(a.chain)]
    x0 = torch.randn((1, a.m), generator=gen, device=dev)
    vecs = [comm.vector(a.m) for _ in range(a.chain)]
    sms = lib.lammcu_sm_count()
    trace = torch.zeros((a.chain * sms * 8,), dtype=torch.int64, device=dev)

    def run():
        for i, w in enumerate(ws):
            comm.mul_mat(t, w, x0 if i == 0 else vecs[i - 1], vecs[i], a.m, a.m)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        run()
        run()
    torch.cuda.synchronize()
    dist.barrier()
    lib.lammcu_set_trace(trace.data_ptr(), trace.numel())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        run()
    lib.lammcu_set_trace(None, 0)
    for _ in range(3):
        dist.barrier()
        g.replay()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    e0.record()
    for _ in range(20):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 20 / a.chain
    tr = trace.cpu().numpy().reshape(a.chain, sms, 8).astype(np.float64)
    cyc = 1e3 / a.mhz
    if rank == 0:
        print(f"{a.fmt} {a.m}x{a.m} rows / {world}, {a.chain} chained sharded launches, {us:.2f} us per launch back to back; "
              f"ns, medians over rank 0's CTAs")
        print("launch  start-prev_start  poll_done  x_quantized  lane_setup  first_weights  row0  cta_done")
        prev = None
        for i in range(a.chain):
            r = tr[i]
            r = r[r[:, 4] > 0]
            if not len(r):
                continue
            start = np.median(r[:, 0])
            rel = lambda e: np.median((r[:, e] - r[:, 1]) * cyc)
            print(f"{i:5d} {(start - prev) if prev else 0:16.0f} {rel(2):10.0f} {rel(3):12.0f} {rel(5):11.0f} {rel(6):14.0f} {rel(7):5.0f} {rel(4):9.0f}")
            prev = start
    assert not comm.timed_out()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
