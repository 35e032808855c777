"""Row sharding of the LAMM mul_mat across the GPUs of one box.

One process per GPU (torch.distributed is the plumbing: rank discovery and the exchange of
64-byte IPC handles).  Weight rows are split contiguously, GPU g owns rows
[g*ceil(M/G), (g+1)*ceil(M/G)) -- the reference's thread split (src/lamm_impl.hpp:107-112) and
upstream's device split (llama.cpp-b2430/ggml-cuda.cu:8480-8492) -- activations are replicated, and
the output slices reach every rank's full C vector through the decode kernel's peer-store epilogue
(lammcu_mul_mat_sharded_dev; {value, tag} pairs in one 64-bit NVLink store, polled by the consumer):
no collective call, fence or flag on the data path.
"""
import ctypes as C


def row_shard(m, rank, world):
    """Rows [lo, hi) of an M-row weight owned by `rank` (contiguous, last shard takes the remainder)."""
    per = (m + world - 1) // world
    lo = min(m, rank * per)
    return lo, min(m, lo + per)


def shard_sizes(m, world):
    return [row_shard(m, g, world)[1] - row_shard(m, g, world)[0] for g in range(world)]


def assemble_rows(slices, m, world):
    """Inverse of the split: per-rank [N, rows_g] outputs -> [N, M] (host-side check of the layout)."""
    import numpy as np
    out = np.zeros((slices[0].shape[0], m), slices[0].dtype)
    for g, s in enumerate(slices):
        lo, hi = row_shard(m, g, world)
        assert s.shape[1] == hi - lo, "slice %d has %d columns, plan says %d" % (g, s.shape[1], hi - lo)
        out[:, lo:hi] = s
    return out


class SharedBuffer:
    """A device buffer of `nbytes` on every rank, each mapped into every process (cudaIpc)."""

    def __init__(self, lamm, dist, nbytes):
        self.lamm, self.nbytes = lamm, nbytes
        L = lamm.lib()
        rank, world = dist.get_rank(), dist.get_world_size()
        ptr, handle = C.c_void_p(), C.create_string_buffer(64)
        lamm.check(L.lammcu_ipc_alloc(nbytes, C.byref(ptr), handle), "ipc_alloc")
        handles = [None] * world
        dist.all_gather_object(handles, handle.raw)
        self.rank, self.world, self.local = rank, world, ptr.value
        self.ptrs = []
        for g in range(world):
            if g == rank:
                self.ptrs.append(ptr.value)
            else:
                p = C.c_void_p()
                lamm.check(L.lammcu_ipc_open(handles[g], C.byref(p)), "ipc_open")
                self.ptrs.append(p.value)
        dist.barrier()

    def ptr_array(self):
        return (C.c_void_p * self.world)(*self.ptrs)

    def to_tensor(self, torch, dtype, numel, stream_ptr=None):
        """Copy of the LOCAL buffer as a torch tensor."""
        t = torch.empty((numel,), dtype=dtype, device="cuda")
        self.lamm.check(self.lamm.lib().lammcu_memcpy_async(t.data_ptr(), self.local, t.numel() * t.element_size(),
                                                            stream_ptr or self.lamm._stream_ptr()), "memcpy")
        return t

    def close(self):
        L = self.lamm.lib()
        for g, p in enumerate(self.ptrs):
            if g != self.rank:
                L.lammcu_ipc_close(p)
        L.lammcu_ipc_free(self.local)


class LLVector(SharedBuffer):
    """An M-element output vector of sharded launches in LL format (lammcu.h): M x {f32 value,
    u32 tag}, one buffer per rank, each mapped into every process."""

    def __init__(self, lamm, dist, m):
        self.m = m
        self.seq = None            # communicator sequence number of the launch that last wrote it
        super().__init__(lamm, dist, lamm.lib().lammcu_ll_bytes(m))


class _SoloDist:
    """torch.distributed stand-in for world_size 1 (a single GPU chaining launches through LL vectors)."""

    def get_rank(self):
        return 0

    def get_world_size(self):
        return 1

    def all_gather_object(self, out, obj):
        out[0] = obj

    def barrier(self):
        pass


class Comm:
    """The C-ABI communicator plus the host-side bookkeeping of launch sequence numbers (the kernel
    derives the same numbers on the device; the host only needs DISTANCES between launches)."""

    def __init__(self, lamm, dist=None):
        self.lamm, self.dist = lamm, dist or _SoloDist()
        self.rank, self.world = self.dist.get_rank(), self.dist.get_world_size()
        h = C.c_void_p()
        lamm.check(lamm.lib().lammcu_comm_create(self.rank, self.world, C.byref(h)), "comm_create")
        self.handle = h
        self.seq = 0               # sharded launches issued (or captured) so far
        self._flags = None         # prefill completion flags (one u32 per rank), allocated on first use

    def vector(self, m):
        return LLVector(self.lamm, self.dist, m)

    def mul_mat(self, w_type, w_shard, x, out, m, k):
        """w_shard: cuda uint8 [rows, row_bytes] = this rank's rows of an M-row weight; x: the
        activation row, either a device pointer (int) / cuda f32 tensor of K plain floats or an
        LLVector produced by an earlier call; out: LLVector of M elements (the full C on every rank)."""
        lo, hi = row_shard(m, self.rank, self.world)
        assert w_shard.shape[0] == hi - lo and out.m >= m
        if isinstance(x, LLVector):
            assert x.seq is not None, "LL activation vector was never produced"
            xp, back = x.local, self.seq - x.seq
            assert back >= 1
        else:
            xp, back = (x if isinstance(x, int) else x.data_ptr()), 0
        self.lamm.check(self.lamm.lib().lammcu_mul_mat_sharded_dev(
            self.handle, w_type, w_shard.data_ptr(), w_shard.stride(0) * w_shard.element_size(), xp, back,
            out.ptr_array(), lo, hi - lo, k, self.lamm._stream_ptr()), "mul_mat_sharded")
        out.seq = self.seq
        self.seq += 1

    def mul_mat_batch(self, w_type, w_shards, x, outs, ms, k):
        """Sibling matrices sharing the activation row x (q/k/v, gate/up) as ONE sharded launch:
        w_shards[j] = this rank's rows of matrix j (ms[j] rows in total), outs[j] its LLVector."""
        n = len(w_shards)
        assert n == len(outs) == len(ms) and 1 <= n <= 4
        if isinstance(x, LLVector):
            assert x.seq is not None, "LL activation vector was never produced"
            xp, back = x.local, self.seq - x.seq
            assert back >= 1
        else:
            xp, back = (x if isinstance(x, int) else x.data_ptr()), 0
        los = [row_shard(m, self.rank, self.world) for m in ms]
        for w, (lo, hi), o, m in zip(w_shards, los, outs, ms):
            assert w.shape[0] == hi - lo and o.m >= m
        stride = w_shards[0].stride(0) * w_shards[0].element_size()
        wp = (C.c_void_p * n)(*[w.data_ptr() for w in w_shards])
        cp = (C.c_void_p * (n * self.world))(*[p for o in outs for p in o.ptrs])
        rb = (C.c_int64 * n)(*[lo for lo, _ in los])
        rows = (C.c_int64 * n)(*[hi - lo for lo, hi in los])
        self.lamm.check(self.lamm.lib().lammcu_mul_mat_sharded_batch_dev(
            self.handle, w_type, n, wp, stride, xp, back, cp, rb, rows, k, self.lamm._stream_ptr()),
            "mul_mat_sharded_batch")
        for o in outs:
            o.seq = self.seq
        self.seq += 1

    def result_buffer(self, n, m):
        """Full f32 prefill result [n][m] (ggml layout), one buffer per rank mapped into every process."""
        return SharedBuffer(self.lamm, self.dist, n * m * 4)

    def mul_mat_gemm(self, w_type, w_shard, x, c_buf, m, n, k):
        """Row-sharded prefill: x = cuda f32 [n, k] (replicated), w_shard = this rank's rows of an M-row
        weight, c_buf = result_buffer(n, m); afterwards (stream order) the LOCAL buffer holds all of C."""
        lo, hi = row_shard(m, self.rank, self.world)
        assert w_shard.shape[0] == hi - lo and x.shape == (n, k)
        if self._flags is None:
            self._flags = SharedBuffer(self.lamm, self.dist, 256)
        fp32 = C.POINTER(C.c_float)
        self.lamm.check(self.lamm.lib().lammcu_mul_mat_sharded_gemm_dev(
            self.handle, w_type, w_shard.data_ptr(), w_shard.stride(0) * w_shard.element_size(),
            C.cast(x.data_ptr(), fp32), x.stride(0), c_buf.ptr_array(), m, self._flags.ptr_array(), lo, hi - lo, n, k,
            self.lamm._stream_ptr()), "mul_mat_sharded_gemm")

    def unpack(self, torch, vec, m=None):
        """LLVector -> plain f32 cuda tensor [m] (waits for every element on the device)."""
        m = m or vec.m
        out = torch.empty((m,), dtype=torch.float32, device="cuda")
        self.lamm.check(self.lamm.lib().lammcu_ll_unpack_dev(self.handle, vec.local, self.seq - vec.seq, out.data_ptr(), m,
                                                             self.lamm._stream_ptr()), "ll_unpack")
        return out

    def sync(self):
        """Wait for the stream; raises LammError if any poll of an earlier sharded call gave up."""
        self.lamm.check(self.lamm.lib().lammcu_comm_sync(self.handle, self.lamm._stream_ptr()), "comm_sync")

    def timed_out(self):
        v = C.c_int(0)
        self.lamm.check(self.lamm.lib().lammcu_comm_status(self.handle, C.byref(v)), "comm_status")
        return bool(v.value)

    def close(self):
        self.lamm.lib().lammcu_comm_destroy(self.handle)
