"""Multi-GPU plumbing: samples-per-pixel shard across ranks, one rank per GPU.
Work unit = (pixel, global spp index); paths never interact and the film is a sum, so the only
exchange step is ONE reduce of the partial films (3*W*H floats) onto rank 0.  The scene/BVH is
replicated.  On the GPU box the reduce is wpt_film_reduce of the C ABI (libwpt.so -> ncclReduce over NVLink, its own
communicator); torch.distributed only ships the 128-byte NCCL id at start-up (FilmReducer) and runs the gloo reduce of
the CPU tests.  The reference has no counterpart (single process, single GPU)."""
import ctypes as C


def spp_range_strong(rank, world, spp_total):
    """Strong scaling: split [0, spp_total) into `world` contiguous, disjoint ranges that differ in
    size by at most one.  Returns (spp_begin, spp_count)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(spp_total, world)
    begin = rank * base + min(rank, rem)
    return begin, base + (1 if rank < rem else 0)


def spp_range_weak(rank, world, spp_per_rank):
    """Weak scaling (bench.py): every rank renders spp_per_rank samples of the same estimator;
    rank r owns global indices [r*spp_per_rank, (r+1)*spp_per_rank)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return rank * spp_per_rank, spp_per_rank


def reduce_film(film, dst=0):
    """Sum the partial radiance buffers onto `dst` (ncclReduce / gloo reduce). No-op when not
    initialised or world == 1.  Returns the total number of ranks."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return 1
    dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM)
    return dist.get_world_size()


class FilmReducer:
    """One NCCL communicator owned by libwpt.so (wpt_nccl_comm_create) + wpt_film_reduce on it.  The rendezvous —
    rank 0's ncclUniqueId to every rank — goes through an already initialised torch.distributed group (any backend);
    the data plane does not: the reduce is enqueued by libwpt on the caller's CUDA stream."""

    def __init__(self, device):
        import torch
        import torch.distributed as dist
        import wpt
        self.L = wpt.lib()
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        idb = (C.c_char * 128)()
        if self.rank == 0:
            wpt._check(self.L.wpt_nccl_unique_id(idb))
        t = torch.frombuffer(bytearray(idb.raw), dtype=torch.uint8).clone()
        if dist.get_backend() == "nccl":
            t = t.cuda(device)
        dist.broadcast(t, src=0)
        raw = bytes(t.cpu().numpy().tobytes())
        self.comm = C.c_void_p()
        wpt._check(self.L.wpt_nccl_comm_create(C.c_char_p(raw), C.c_int(self.world), C.c_int(self.rank), C.c_int(device),
                                               C.byref(self.comm)))

    def reduce(self, film, root=0, stream=None):
        """film: CUDA float32 [3, N] planar; summed in place onto `root`, asynchronously on `stream`"""
        import wpt
        N = film.numel() // 3
        base = film.data_ptr()
        wpt._check(self.L.wpt_film_reduce(self.comm, C.c_void_p(base), C.c_void_p(base + 4 * N), C.c_void_p(base + 8 * N),
                                          C.c_uint32(N), C.c_int(root), wpt._stream_handle(stream)))

    def shard_layout(self, n_pixels):
        """(padded_floats, first, count) of this rank's slice of the flattened [3, N] film (wpt_film_shard_layout)"""
        import wpt
        pad, lo, cnt = C.c_uint64(), C.c_uint64(), C.c_uint64()
        wpt._check(self.L.wpt_film_shard_layout(C.c_uint32(n_pixels), C.c_int(self.rank), C.c_int(self.world),
                                                C.byref(pad), C.byref(lo), C.byref(cnt)))
        return pad.value, lo.value, cnt.value

    def reduce_scatter(self, film, n_pixels, stream=None):
        """film: flat CUDA float32 buffer of at least shard_layout()[0] floats holding the planar [3, N] film; afterwards
        this rank's slice holds the sum over the ranks (wpt_film_reduce_scatter: ncclReduceScatter, in place)"""
        import wpt
        wpt._check(self.L.wpt_film_reduce_scatter(self.comm, C.c_void_p(film.data_ptr()), C.c_uint32(n_pixels),
                                                  C.c_int(self.rank), C.c_int(self.world), wpt._stream_handle(stream)))

    def shard_finish(self, film, n_pixels, scale, host_film=None, stream=None):
        """scale this rank's slice and queue its copy into host_film (numpy float32 over pinned / registered memory of
        3 * N floats: the whole film, every rank writes its own part)"""
        import wpt
        hp = C.c_void_p(host_film.ctypes.data) if host_film is not None else None
        wpt._check(self.L.wpt_film_shard_finish(C.c_void_p(film.data_ptr()), C.c_uint32(n_pixels), C.c_int(self.rank),
                                                C.c_int(self.world), C.c_float(scale), hp, wpt._stream_handle(stream)))

    def close(self):
        if self.comm:
            self.L.wpt_nccl_comm_destroy(self.comm)
            self.comm = C.c_void_p()


class SharedHostFilm:
    """The host film of a frame rendered by one process per GPU: a file in `directory` (a tmpfs such as /dev/shm) mapped
    by every rank and page-locked in each (wpt_host_register), so that every rank's slice goes to its place by DMA.
    `array` is the flat float32 [3 * N] film.  Rank 0 creates the file; call after the ranks agreed on `directory`."""

    def __init__(self, directory, n_floats, rank, barrier):
        import os
        import numpy as np
        import wpt
        self.path = os.path.join(directory, "film.f32")
        nbytes = (4 * n_floats + 4095) // 4096 * 4096          # whole pages: that is what gets page-locked
        if rank == 0:
            with open(self.path, "wb") as f:
                f.truncate(nbytes)
        barrier()
        self._map = np.memmap(self.path, dtype=np.float32, mode="r+", shape=(nbytes // 4,))
        self.array = self._map[:n_floats]
        self._ptr = C.c_void_p(self._map.ctypes.data)
        wpt._check(wpt.lib().wpt_host_register(self._ptr, C.c_size_t(nbytes)))

    def close(self):
        import wpt
        if self._ptr:
            wpt.lib().wpt_host_unregister(self._ptr)
            self._ptr = None
