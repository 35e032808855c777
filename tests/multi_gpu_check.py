"""Row-sharded decode across the GPUs of one box, checked against the oracle.
Run under torchrun (one rank per GPU):  torchrun --nproc-per-node 2 tests/multi_gpu_check.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def row_empty_somewhere(m, world):
    per = (m + world - 1) // world
    return per * (world - 1) >= m


def main():
    import torch
    import torch.distributed as dist
    import lamm_b200 as L
    from lamm_b200 import sharding
    from oracle_lib import Oracle, Q4_1, Q8_0, Q2_K, Q5_0
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L.init(local)
    comm = sharding.Comm(L, dist)
    o = Oracle()
    ok = True
    for t, m, k in [(Q4_1, 4096, 4096), (Q8_0, 1000, 2048), (Q5_0, 11008, 4096), (Q2_K, 515, 11008)]:
        rng = np.random.default_rng(7)                      # same data on every rank
        w = o.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=3)
        x = rng.standard_normal((1, k)).astype(np.float32)
        want = o.mul_mat(t, w, x)
        lo, hi = sharding.row_shard(m, rank, world)
        wd = torch.from_numpy(w[lo:hi].copy()).cuda()
        xd = torch.from_numpy(x).cuda()
        cbuf = sharding.LLVector(L, dist, m)
        for it in range(3):                                 # repeated launches exercise the tag protocol
            comm.mul_mat(t, wd, xd, cbuf, m, k)
        # a follow-up sharded launch that READS the gathered vector (every rank needs all of it)
        w2 = o.quantize_weights(Q4_1, np.random.default_rng(9).standard_normal((256, (m // 256) * 256)).astype(np.float32))
        k2 = (m // 256) * 256
        lo2, hi2 = sharding.row_shard(256, rank, world)
        c2 = sharding.LLVector(L, dist, 256)
        comm.mul_mat(Q4_1, torch.from_numpy(w2[lo2:hi2].copy()).cuda(), cbuf, c2, 256, k2)
        got_t = comm.unpack(torch, cbuf)
        got2_t = comm.unpack(torch, c2)
        torch.cuda.synchronize()
        dist.barrier()
        got = got_t.cpu().numpy()[None, :]
        tol = 1e-5 * np.sqrt(k) * np.abs(want).max()
        err = np.abs(got - want).max()
        want2 = o.mul_mat(Q4_1, w2, got[:, :k2])
        got2 = got2_t.cpu().numpy()[None, :]
        err2 = np.abs(got2 - want2).max()
        tol2 = 1e-5 * np.sqrt(k2) * np.abs(want2).max()
        good = err <= tol and err2 <= tol2
        ok &= bool(good)
        print(f"rank {rank} type {t} {m}x{k}: err {err:.3g} (tol {tol:.3g}), chained err {err2:.3g} (tol {tol2:.3g}) "
              f"{'ok' if good else 'FAIL'}", flush=True)
        dist.barrier()
    # sibling batch: three matrices sharing x in ONE sharded launch, then a launch that consumes the middle one
    for t, ms, k in [(Q4_1, (4096, 4096, 4096), 4096), (Q5_0, (1024, 512, 2048), 2048), (Q2_K, (512, 1024), 1024)]:
        rng = np.random.default_rng(11)
        ws = [o.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=5 + j) for j, m in enumerate(ms)]
        x = rng.standard_normal((1, k)).astype(np.float32)
        xd = torch.from_numpy(x).cuda()
        shards, outs = [], []
        for w, m in zip(ws, ms):
            lo, hi = sharding.row_shard(m, rank, world)
            shards.append(torch.from_numpy(w[lo:hi].copy()).cuda())
            outs.append(sharding.LLVector(L, dist, m))
        for it in range(2):
            comm.mul_mat_batch(t, shards, xd, outs, list(ms), k)
        k2 = ms[1]
        w2 = o.quantize_weights(Q4_1, np.random.default_rng(13).standard_normal((512, k2)).astype(np.float32))
        lo2, hi2 = sharding.row_shard(512, rank, world)
        c2 = sharding.LLVector(L, dist, 512)
        comm.mul_mat(Q4_1, torch.from_numpy(w2[lo2:hi2].copy()).cuda(), outs[1], c2, 512, k2)
        gots = [comm.unpack(torch, v) for v in outs]
        got2_t = comm.unpack(torch, c2)
        torch.cuda.synchronize()
        dist.barrier()
        good = True
        for w, g in zip(ws, gots):
            want = o.mul_mat(t, w, x)
            good &= bool(np.abs(g.cpu().numpy()[None, :] - want).max() <= 1e-5 * np.sqrt(k) * np.abs(want).max())
        want2 = o.mul_mat(Q4_1, w2, gots[1].cpu().numpy()[None, :])
        good &= bool(np.abs(got2_t.cpu().numpy()[None, :] - want2).max() <= 1e-5 * np.sqrt(k2) * np.abs(want2).max())
        ok &= good
        print(f"rank {rank} type {t} sibling batch {ms}x{k}: {'ok' if good else 'FAIL'}", flush=True)
        dist.barrier()
    # row-sharded PREFILL: every rank ends up with the full [N][M] result through the GEMM's peer-store epilogue
    from oracle_lib import Q5_1, F32
    for t, m, k, n in [(Q4_1, 1024, 1024, 96), (Q5_1, 1000, 512, 200), (Q8_0, 2048, 2048, 130), (Q2_K, 768, 1024, 64),
                       (Q4_1, 512, 256, 5)]:
        rng = np.random.default_rng(21)
        w = o.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=4)
        x = rng.standard_normal((n, k)).astype(np.float32)
        want = o.mul_mat(t, w, x)
        lo, hi = sharding.row_shard(m, rank, world)
        wd = torch.from_numpy(w[lo:hi].copy()).cuda()
        xd = torch.from_numpy(x).cuda()
        cbuf = comm.result_buffer(n, m)
        for it in range(2):
            comm.mul_mat_gemm(t, wd, xd, cbuf, m, n, k)
        got = cbuf.to_tensor(torch, torch.float32, n * m).view(n, m)
        torch.cuda.synchronize()
        dist.barrier()
        err = np.abs(got.cpu().numpy() - want).max()
        tol = 1e-5 * np.sqrt(k) * np.abs(want).max()
        good = bool(err <= tol)
        ok &= good
        print(f"rank {rank} type {t} sharded prefill {m}x{k} N={n}: err {err:.3g} (tol {tol:.3g}) {'ok' if good else 'FAIL'}", flush=True)
        dist.barrier()
    # SURVEY 8f-4, sharded load: every rank opens the committed GGUF fixture, registers ONLY its row range of each 2-D
    # tensor from the file mapping (gguf_init_from_file + the loader's reads, ggml.c:20655), and the sharded decode on
    # those device rows reproduces the oracle on the whole tensor
    from test_gguf import FIXTURE

    class DevRows:                                           # what Comm.mul_mat needs of a cuda uint8 [rows, row_bytes]
        def __init__(self, h, row_bytes):
            self.h, self.rb = h, row_bytes
            self.shape = (int(L.lib().lammcu_weight_rows(h)), row_bytes)

        def data_ptr(self):
            return int(L.lib().lammcu_weight_dev_ptr(self.h))

        def stride(self, d):
            return self.rb

        def element_size(self):
            return 1
    # the committed fixture (written by the reference's gguf writer) has at most 33 rows per tensor; a second file with
    # model-sized row counts is assembled here by rank 0 (version 3, default alignment 32, no key-value pairs)
    import struct
    import tempfile
    big = os.path.join(tempfile.gettempdir(), f"lamm_shard_{os.environ.get('MASTER_PORT', '0')}.gguf")
    big_tensors = [("w.q4_1", Q4_1, 1000, 4096), ("w.q5_0", Q5_0, 515, 2048), ("w.q2_k", Q2_K, 300, 1024), ("w.f16", 1, 64, 512)]
    if rank == 0:
        meta = b"GGUF" + struct.pack("<IQQ", 3, len(big_tensors), 0)
        blobs, off = [], 0
        for name, t, m, k in big_tensors:
            rng = np.random.default_rng(m)
            if t == 1:
                wq = rng.standard_normal((m, k)).astype(np.float16).view(np.uint8).reshape(m, -1)
            else:
                wq = o.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=m)
            meta += struct.pack("<Q", len(name)) + name.encode() + struct.pack("<IQQIQ", 2, k, m, t, off)
            b = wq.tobytes() + b"\0" * ((-wq.nbytes) % 32)
            blobs.append(b)
            off += len(b)
        with open(big, "wb") as f:
            f.write(meta + b"\0" * ((-len(meta)) % 32) + b"".join(blobs))
    dist.barrier()
    for path, names in [(FIXTURE, ("blk.0.attn_q.weight", "blk.0.attn_k.weight", "blk.0.ffn_down.weight",
                                   "blk.0.ffn_up.weight", "output.weight")),
                        (big, tuple(n for n, _, _, _ in big_tensors))]:
      with L.Gguf(path) as g:
        try:
            for name in names:
                i = g.find(name)
                if row_empty_somewhere(g.tensor(i)[3][1], world):
                    continue                                     # fewer rows than ranks: not a sharding case
                _, t, _, ne, raw = g.tensor(i)
                k, m = ne[0], ne[1]
                wq = raw.reshape(m, -1)
                x = np.random.default_rng(m + k).standard_normal((1, k)).astype(np.float32)
                want = o.mul_mat(t, wq, x)
                lo, hi = sharding.row_shard(m, rank, world)
                good = True
                h = g.upload_rows(i, lo, hi)
                good &= int(L.lib().lammcu_weight_bytes(h)) == (hi - lo) * wq.shape[1]         # its byte range, no more
                shard = DevRows(h, wq.shape[1])
                if t == 1:                                       # F16 rows: the row-sharded prefill GEMM (N = 40)
                    x = np.random.default_rng(m + k).standard_normal((40, k)).astype(np.float32)
                    want = o.mul_mat(t, wq, x)
                    cb = comm.result_buffer(40, m)
                    comm.mul_mat_gemm(t, shard, torch.from_numpy(x).cuda(), cb, m, 40, k)
                    got_t = cb.to_tensor(torch, torch.float32, 40 * m).view(40, m)
                    torch.cuda.synchronize()
                    dist.barrier()
                    err = np.abs(got_t.cpu().numpy() - want).max()
                else:
                    cbuf = sharding.LLVector(L, dist, m)
                    comm.mul_mat(t, shard, torch.from_numpy(x).cuda(), cbuf, m, k)
                    got_t = comm.unpack(torch, cbuf)
                    torch.cuda.synchronize()
                    dist.barrier()
                    err = np.abs(got_t.cpu().numpy()[None, :] - want).max()
                good &= bool(err <= 1e-5 * np.sqrt(k) * np.abs(want).max())
                ok &= good
                print(f"rank {rank} gguf {name} rows {lo}:{hi} of {m}: err {err:.3g} {'ok' if good else 'FAIL'}", flush=True)
                dist.barrier()
        finally:
            torch.cuda.synchronize()
            L.lib().lammcu_weight_forget_all()
    dist.barrier()
    if rank == 0:
        os.unlink(big)
    comm.sync()                                              # raises if any LL poll gave up
    ok &= not comm.timed_out()
    flag = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(flag)
    dist.barrier()
    dist.destroy_process_group()
    if flag.item():
        sys.exit(1)
    if rank == 0:
        print("multi_gpu_check: all ok")


if __name__ == "__main__":
    main()
