"""the GGUF reader (csrc/lamm_gguf.cpp, include/lammcu.h lammcu_gguf_*) against files written by
the REFERENCE's own writer (gguf_add_tensor + gguf_write_to_file through oracle/_ref/libref.so), and against a
committed fixture of the same kind for machines without the compiled reference.  CPU only: parsing needs no
device.  Robustness: every truncation of a valid file and single-byte corruptions of its header must be an
error or a consistent parse, never a crash."""
import os
import sys

import numpy as np
import pytest
from conftest import GOLDEN
from oracle_lib import F16, F32, Q2_K, Q4_K, Q6_K, row_size

FIXTURE = os.path.join(GOLDEN, "gguf_fixture_v3.gguf")
I32 = 26


def fixture_tensors(oracle):
    """(name, type, bytes [ne1, row bytes], ne0, ne1); ne1 == 0: 1-D.  Deterministic."""
    rng = np.random.default_rng(77)
    out = []
    for name, t, ne0, ne1 in [("token_embd.weight", 3, 64, 10), ("blk.0.attn_q.weight", 3, 256, 33),
                              ("blk.0.attn_k.weight", 8, 256, 7), ("blk.0.ffn_down.weight", Q2_K, 512, 5),
                              ("blk.0.ffn_up.weight", Q4_K, 256, 9), ("blk.0.ffn_gate.weight", Q6_K, 256, 3),
                              ("blk.0.attn_norm.weight", F32, 256, 0), ("output.weight", F16, 96, 4),
                              ("ids", I32, 12, 0)]:
        rows = max(ne1, 1)
        if t == I32:
            raw = rng.integers(0, 1000, (rows, ne0)).astype(np.int32).view(np.uint8)
        elif t == F32:
            raw = rng.standard_normal((rows, ne0)).astype(np.float32).view(np.uint8)
        else:
            raw = oracle.quantize_weights(t, rng.standard_normal((rows, ne0)).astype(np.float32), seed=ne0 + ne1)
        out.append((name, t, np.ascontiguousarray(raw).reshape(rows, -1), ne0, ne1))
    return out


def check_against(lamm, path, want, alignment):
    with lamm.Gguf(path) as g:
        assert len(g) == len(want) and g.alignment == alignment and g.data_offset % alignment == 0
        assert g.get_str("general.architecture") == "llama" and g.get_str("general.name") == "lamm gguf fixture"
        assert g.get_u32("llama.block_count") == 32 and abs(g.get_f32("llama.rope.freq_base") - 10000.5) < 1e-3
        assert g.get_u32("no.such.key") is None and g.get_u32("general.name") is None       # wrong kind -> absent
        assert g.get_str("llama.block_count") is None
        for i, (name, t, raw, ne0, ne1) in enumerate(want):
            got_name, got_t, nd, ne, data = g.tensor(i)
            assert (got_name, got_t) == (name, t)
            assert nd == (2 if ne1 else 1) and ne == [ne0, ne1 if ne1 else 1, 1, 1]
            assert np.array_equal(data, raw.reshape(-1)), name
            assert g.find(name) == i
        assert g.find("missing") == -1
        with pytest.raises(lamm.LammError):
            g.tensor(len(want))


def test_reader_vs_reference_writer(lamm, oracle, ref, tmp_path):
    """Default alignment only: the reference's writer pads with the context's alignment, which
    gguf_set_val_u32("general.alignment") does not change (ggml.c:21175-21200), so a file it writes with that
    key set is not self-consistent.  Other alignments are covered by test_alignment_key_is_honoured."""
    want = fixture_tensors(oracle)
    path = str(tmp_path / "ref.gguf")
    ref.gguf_write(path, want)
    check_against(lamm, path, want, 32)
    assert open(path, "rb").read() == open(FIXTURE, "rb").read()          # the committed fixture is this very file


def test_alignment_key_is_honoured(lamm, tmp_path):
    """A hand-assembled version-3 file with general.alignment = 256 and two tensors."""
    import struct

    def s(b):
        return struct.pack("<Q", len(b)) + b
    t0 = np.arange(64, dtype=np.float32)
    t1 = np.arange(20 * 2, dtype=np.uint8)                     # Q4_1, K = 32, 2 rows
    meta = b"GGUF" + struct.pack("<IQQ", 3, 2, 1) + s(b"general.alignment") + struct.pack("<II", 4, 256)
    meta += s(b"a") + struct.pack("<IQIQ", 1, 64, 0, 0)
    meta += s(b"b") + struct.pack("<IQQIQ", 2, 32, 2, 3, 256)
    pad = (-len(meta)) % 256
    blob = meta + b"\0" * pad + t0.tobytes() + t1.tobytes()
    path = str(tmp_path / "al.gguf")
    open(path, "wb").write(blob)
    with lamm.Gguf(path) as g:
        assert g.alignment == 256 and g.data_offset == len(meta) + pad and len(g) == 2
        assert np.array_equal(g.tensor(0)[4].view(np.float32), t0)
        name, t, nd, ne, raw = g.tensor(1)
        assert (name, t, nd, ne) == ("b", 3, 2, [32, 2, 1, 1]) and np.array_equal(raw, t1)
    open(path, "wb").write(blob.replace(struct.pack("<II", 4, 256), struct.pack("<II", 4, 96), 1))
    with pytest.raises(lamm.LammError):                         # not a power of two
        lamm.Gguf(path)
    open(path, "wb").write(meta[:-8] + struct.pack("<Q", 8) + b"\0" * pad + t0.tobytes() + t1.tobytes())
    with pytest.raises(lamm.LammError):                         # tensor offset not aligned
        lamm.Gguf(path)


def test_committed_fixture(lamm, oracle):
    """The same file, written once by the reference here (oracle/gen_golden.py) and committed."""
    check_against(lamm, FIXTURE, fixture_tensors(oracle), 32)
    for t, k in [(3, 256), (Q2_K, 512)]:                       # the sizes the fixture relies on
        assert row_size(t, k) == lamm.row_size(t, k)


def test_truncated_and_corrupt_files_fail_cleanly(lamm, tmp_path):
    blob = open(FIXTURE, "rb").read()
    with lamm.Gguf(FIXTURE) as g:
        data_offset, n = g.data_offset, len(g)
    path = str(tmp_path / "bad.gguf")
    # every truncation inside the metadata, and a sample inside the data section, must be refused
    cuts = list(range(0, data_offset + 1, 7)) + list(range(data_offset + 1, len(blob) - 1, 997))
    for cut in cuts:
        open(path, "wb").write(blob[:cut])
        with pytest.raises(lamm.LammError):
            lamm.Gguf(path)
    # single-byte damage in the metadata: either refused, or parsed into in-bounds tensors
    rng = np.random.default_rng(1)
    for pos in rng.integers(0, data_offset, 300):
        b = bytearray(blob)
        b[pos] ^= 0xFF
        open(path, "wb").write(bytes(b))
        try:
            g = lamm.Gguf(path)
        except lamm.LammError:
            continue
        with g:
            for i in range(len(g)):
                g.tensor(i)                                   # copies the bytes: would fault if out of bounds
    open(path, "wb").write(b"GGML" + blob[4:])
    with pytest.raises(lamm.LammError):
        lamm.Gguf(path)
    with pytest.raises(lamm.LammError):
        lamm.Gguf(str(tmp_path / "does_not_exist.gguf"))
    assert n == 9


def test_hostile_dimensions_cannot_wrap_the_size_check(lamm, tmp_path):
    """Dimensions whose byte count wraps 64 (or 128) bits must not pass as 'inside the file': F32 ne0 = 2^62 + 8 has a
    row size of 32 bytes modulo 2^64; 2^32 x 2^32 x 2^32 x 2^32 rows of 4 bytes is 0 modulo 2^128."""
    import struct

    def s(b):
        return struct.pack("<Q", len(b)) + b
    path = str(tmp_path / "hostile.gguf")
    for dims, typ in [((2 ** 62 + 8, 1), 0), ((2 ** 32, 2 ** 32, 2 ** 32, 2 ** 32), 0), ((32, 2 ** 63 - 1), 3),
                      ((2 ** 62, 4), 1), ((2 ** 63 - 256, 1), 10), ((8, 2 ** 61, 8), 0)]:
        meta = b"GGUF" + struct.pack("<IQQ", 3, 1, 0) + s(b"t") + struct.pack("<I", len(dims))
        meta += b"".join(struct.pack("<Q", d) for d in dims) + struct.pack("<IQ", typ, 0)
        open(path, "wb").write(meta + b"\0" * ((-len(meta)) % 32) + b"\0" * 64)
        with pytest.raises(lamm.LammError):
            lamm.Gguf(path)
    meta = b"GGUF" + struct.pack("<IQQ", 3, 1, 0) + s(b"t") + struct.pack("<IQQIQ", 2, 8, 2, 0, 0)     # 8 x 2 F32: fits
    open(path, "wb").write(meta + b"\0" * ((-len(meta)) % 32) + b"\0" * 64)
    with lamm.Gguf(path) as g:
        assert g.tensor(0)[3] == [8, 2, 1, 1] and g.tensor(0)[4].nbytes == 64


def test_row_ranges_are_contiguous_byte_ranges(lamm):
    with lamm.Gguf(FIXTURE) as g:
        for i in range(len(g)):
            _, t, nd, ne, raw = g.tensor(i)
            m = ne[1]
            rb = len(raw) // m
            for lo, hi in [(0, m), (0, 1), (m - 1, m), (m // 3, m // 3 + max(1, m // 2))]:
                hi = min(hi, m)
                assert np.array_equal(g.rows(i, lo, hi), raw[lo * rb:hi * rb])
            for lo, hi in [(-1, 1), (0, m + 1), (1, 1)]:
                with pytest.raises(lamm.LammError):
                    g.rows(i, lo, hi)


def test_rank_shards_tile_every_tensor(lamm):
    """SURVEY 8f-4, host half of the sharded load: the byte ranges the ranks of a 2-, 3- and 8-way row shard would upload
    (lammcu_gguf_rows on sharding.row_shard) are disjoint, ordered and concatenate to the tensor's bytes."""
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from lamm_b200.sharding import row_shard
    with lamm.Gguf(FIXTURE) as g:
        for i in range(len(g)):
            _, t, nd, ne, raw = g.tensor(i)
            m = ne[1]
            for world in (2, 3, 8):
                parts = []
                for rank in range(world):
                    lo, hi = row_shard(m, rank, world)
                    if hi > lo:
                        parts.append(g.rows(i, lo, hi))
                assert np.array_equal(np.concatenate(parts), raw), (i, world)


def test_upload_needs_a_device_and_a_weight_type(lamm):
    """lammcu_gguf_upload_rows: refused for 1-D / unsupported-type tensors and bad row ranges; without a CUDA
    device the registration itself fails loudly (no CPU fallback)."""
    with lamm.Gguf(FIXTURE) as g:
        for name in ("ids", "output.weight"):                  # I32, F16: not weight types of the path
            i = g.find(name)
            with pytest.raises(lamm.LammError):
                g.upload_rows(i, 0, g.tensor(i)[3][1])
        i = g.find("blk.0.attn_q.weight")
        for lo, hi in [(-1, 3), (0, 34), (5, 5), (7, 3)]:
            with pytest.raises(lamm.LammError):
                g.upload_rows(i, lo, hi)
        import torch
        if not torch.cuda.is_available():
            with pytest.raises(lamm.LammError):
                g.upload_rows(i, 0, 33)
