"""Row sharding of the LAMM mul_mat across the GPUs of one box.

One process per GPU (torch.distributed is the plumbing: rank discovery and the exchange of
64-byte IPC handles).  Weight rows are split contiguously, GPU g owns rows
[g*ceil(M/G), (g+1)*ceil(M/G)) -- the reference's thread split (src/lamm_impl.hpp:107-112) and
upstream's device split (llama.cpp-b2430/ggml-cuda.cu:8480-8492) -- activations are replicated, and
the output slices reach every rank's full C vector through the decode kernel's peer-store epilogue
(lammcu_mul_mat_sharded_dev): no collective call on the data path.
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


class Comm:
    """The flag arrays + the C-ABI communicator."""

    def __init__(self, lamm, dist):
        self.lamm, self.dist = lamm, dist
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.flags = SharedBuffer(lamm, dist, 256)
        h = C.c_void_p()
        lamm.check(lamm.lib().lammcu_comm_create(self.rank, self.world, self.flags.ptr_array(), C.byref(h)), "comm_create")
        self.handle = h

    def mul_mat(self, w_type, w_shard, x_ptr, c_buf, m, k):
        """w_shard: cuda uint8 [rows, row_bytes] = this rank's rows of an M-row weight; x_ptr: device
        pointer (int) of the f32 activation row; c_buf: SharedBuffer of >= 4*M bytes (the full C)."""
        lo, hi = row_shard(m, self.rank, self.world)
        assert w_shard.shape[0] == hi - lo
        self.lamm.check(self.lamm.lib().lammcu_mul_mat_sharded_dev(
            self.handle, w_type, w_shard.data_ptr(), w_shard.stride(0) * w_shard.element_size(), x_ptr,
            c_buf.ptr_array(), lo, hi - lo, k, self.lamm._stream_ptr()), "mul_mat_sharded")

    def close(self):
        self.lamm.lib().lammcu_comm_destroy(self.handle)
        self.flags.close()
