"""Host-side logic of the row-sharded path, on CPU: the shard plan and, with two gloo
ranks, that per-rank row slices of the oracle's result assemble to the oracle's full result.  The
CUDA peer-store path itself is tests/multi_gpu_check.py (torchrun, >= 2 GPUs)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def test_row_shard_covers_every_row_once():
    from lamm_b200.sharding import row_shard, shard_sizes
    for m in (1, 7, 37, 256, 4096, 11008, 32000, 17920):
        for world in (1, 2, 3, 4, 8):
            seen = np.zeros(m, np.int32)
            prev_hi = 0
            for g in range(world):
                lo, hi = row_shard(m, g, world)
                assert lo == prev_hi or lo == hi == m, "shards are contiguous and ordered"
                assert 0 <= lo <= hi <= m
                seen[lo:hi] += 1
                prev_hi = max(prev_hi, hi)
            assert (seen == 1).all(), f"M={m} world={world}: every row owned exactly once (no dropped remainder, SURVEY 9-3)"
            assert sum(shard_sizes(m, world)) == m


def test_row_shard_llama_shapes_are_even():
    from lamm_b200.sharding import shard_sizes
    for m in (4096, 11008, 32000, 5120, 13824, 6656, 17920):
        sizes = shard_sizes(m, 8)
        assert len(set(sizes)) == 1, (m, sizes)


def test_assemble_rows_rejects_wrong_slice():
    from lamm_b200.sharding import assemble_rows
    with pytest.raises(AssertionError):
        assemble_rows([np.zeros((1, 3), np.float32), np.zeros((1, 3), np.float32)], 7, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from lamm_b200.sharding import assemble_rows, row_shard
    from oracle_lib import Oracle, Q2_K, Q4_1, Q5_0
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    o = Oracle()
    ok = True
    for t, m, n, k in [(Q4_1, 37, 1, 512), (Q5_0, 129, 3, 256), (Q2_K, 33, 2, 512)]:
        rng = np.random.default_rng(11)                       # same tensors on both ranks
        w = o.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=5)
        x = rng.standard_normal((n, k)).astype(np.float32)
        lo, hi = row_shard(m, rank, world)
        mine = o.mul_mat(t, w[lo:hi], x)                      # [n, rows]: this rank's columns of C
        per = (m + world - 1) // world
        pad = np.zeros((n, per), np.float32)
        pad[:, :hi - lo] = mine
        bufs = [torch.zeros((n, per)) for _ in range(world)]
        dist.all_gather(bufs, torch.from_numpy(pad))          # the exchange step of the path
        slices = [b.numpy()[:, :row_shard(m, g, world)[1] - row_shard(m, g, world)[0]] for g, b in enumerate(bufs)]
        full = assemble_rows(slices, m, world)
        ok &= bool(np.array_equal(full, o.mul_mat(t, w, x)))  # rows are independent: bit-identical
    # each rank takes its row shard straight from the GGUF file (one contiguous byte range of the
    # mapping); the per-rank results assemble to the whole tensor's result
    import lamm_b200
    from test_gguf import FIXTURE
    with lamm_b200.Gguf(FIXTURE) as g:
        for name in ("blk.0.attn_q.weight", "blk.0.ffn_down.weight", "blk.0.ffn_up.weight"):
            i = g.find(name)
            _, t, _, ne, raw = g.tensor(i)
            k, m = ne[0], ne[1]
            x = np.random.default_rng(3).standard_normal((2, k)).astype(np.float32)
            lo, hi = row_shard(m, rank, world)
            shard = g.rows(i, lo, hi).reshape(hi - lo, -1)
            ok &= bool(np.array_equal(shard, raw.reshape(m, -1)[lo:hi]))
            per = (m + world - 1) // world
            pad = np.zeros((2, per), np.float32)
            pad[:, :hi - lo] = o.mul_mat(t, shard, x)
            bufs = [torch.zeros((2, per)) for _ in range(world)]
            dist.all_gather(bufs, torch.from_numpy(pad))
            slices = [b.numpy()[:, :row_shard(m, r, world)[1] - row_shard(m, r, world)[0]] for r, b in enumerate(bufs)]
            ok &= bool(np.array_equal(assemble_rows(slices, m, world), o.mul_mat(t, raw.reshape(m, -1), x)))
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, ok))


def test_two_rank_gloo_assembly_matches_oracle():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True), (1, True)]
