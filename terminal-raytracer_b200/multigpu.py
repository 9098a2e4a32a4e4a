"""
Row-sharded multi-GPU frames: one process per GPU (torch.distributed), one trt context per process.

The path shards by pixels with no exchange while tracing (every pixel depends only on (x, y, uniforms,
scene): shader.wgsl:499-621).  Rows are dealt to ranks in stripes (trt_config.shard_*); RNG seeds and
pixel indices use global coordinates, so the union of the shards is bit-identical to a 1-GPU frame.

Per-frame combine = ROW-ADDRESSED STREAMS: a frame's byte stream is its rows back to back
(lib.rs:503-532) and a row's bytes depend on that row only, so no pixel has to move.  Every rank
quantises its own rows on its GPU (trt_quantise_rows_async), the ranks exchange the byte length of
every row with ONE NCCL all-gather (4 B per row: 8.6 KB for a 2160-row frame, against 132.7 MB of
float4 pixels), every rank adds the lengths up to the offsets of its rows and copies them device ->
host over ITS OWN PCIe link, straight to their place in one host buffer all ranks of the node share
(a POSIX shared-memory segment, page-locked in every process).  NCCL is used for that combine step
only (north_star).  The round-1 combine -- an NCCL gather of the packed float4 rows to rank 0, which
then quantises the assembled frame and copies 196 MB over a single link -- is kept as
`gather_pixels()` for callers that want the float frame on one GPU.

`shard_rows` / `assemble_rows` / `stream_row_offsets` / `assemble_stream` are the host-side statement
of the layouts, used by the gloo tests.
"""
from __future__ import annotations

import ctypes as C

import numpy as np


def shard_rows(height: int, shard_index: int, shard_count: int, stripe_rows: int) -> np.ndarray:
    """Rows owned by a shard: stripes of `stripe_rows` rows dealt round-robin (same rule as
    trt_shard_rows in csrc/trt_api.cu)."""
    y = np.arange(height, dtype=np.uint32)
    if shard_count <= 1:
        return y
    return y[(y // max(1, stripe_rows)) % shard_count == shard_index]


def max_rows_per_shard(height: int, shard_count: int, stripe_rows: int) -> int:
    return max(len(shard_rows(height, s, shard_count, stripe_rows)) for s in range(max(1, shard_count)))


def assemble_rows(blocks, width: int, height: int, shard_count: int, stripe_rows: int) -> np.ndarray:
    """blocks[r]: (>= rows_r * width, 4) float32 packed rows of shard r (local row order).
    Returns the (height * width, 4) frame -- what trt_quantise_gathered reads through its row map."""
    out = np.zeros((height, width, 4), np.float32)
    for r in range(shard_count):
        rows = shard_rows(height, r, shard_count, stripe_rows)
        b = np.asarray(blocks[r], np.float32).reshape(-1, 4)[: len(rows) * width].reshape(len(rows), width, 4)
        out[rows] = b
    return out.reshape(height * width, 4)


def stream_row_offsets(lens_by_rank, height: int, shard_count: int, stripe_rows: int) -> np.ndarray:
    """lens_by_rank[r][k] = byte length of shard r's k-th row (what the all-gather delivers, padded to the
    longest shard).  Returns height + 1 uint64 offsets of the whole frame's rows; [-1] = stream length."""
    lens = np.zeros(height, np.uint64)
    for r in range(shard_count):
        rows = shard_rows(height, r, shard_count, stripe_rows)
        lens[rows] = np.asarray(lens_by_rank[r]).reshape(-1)[: len(rows)].astype(np.uint64)
    off = np.zeros(height + 1, np.uint64)
    np.cumsum(lens, out=off[1:])
    return off


def assemble_stream(streams, lens_by_rank, height: int, shard_count: int, stripe_rows: int) -> bytes:
    """Host statement of trt_read_stream_rows over all shards: streams[r] = shard r's rows back to back."""
    off = stream_row_offsets(lens_by_rank, height, shard_count, stripe_rows)
    out = bytearray(int(off[-1]))
    for r in range(shard_count):
        rows = shard_rows(height, r, shard_count, stripe_rows)
        pos = 0
        for k, y in enumerate(rows):
            n = int(np.asarray(lens_by_rank[r]).reshape(-1)[k])
            out[int(off[y]): int(off[y]) + n] = streams[r][pos: pos + n]
            pos += n
    return bytes(out)


class _DevArray:
    """a raw device pointer as something torch.as_tensor can wrap without a copy"""

    def __init__(self, ptr, n, typestr="<i4"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


class ShardedRenderer:
    """One rank's part of a row-sharded frame.  Needs torch.distributed initialised (NCCL on GPUs)."""

    def __init__(self, pkg, width, height, rank, world, stripe_rows=1, device=0, conv=0):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.pkg = torch, dist, pkg
        self.w, self.h, self.rank, self.world, self.stripe = width, height, rank, world, stripe_rows
        self.r = pkg.Renderer(width, height, device=device, conv=conv, shard_index=rank, shard_count=world,
                              stripe_rows=stripe_rows)
        self.stream = torch.cuda.ExternalStream(self.r.cuda_stream(), device=device)
        self.max_rows = max_rows_per_shard(height, world, stripe_rows)
        self.dev = torch.device("cuda", device)
        # the context's row-length buffer (u32 per owned row) seen as a torch tensor, and the all-gathered table
        ptr, nbytes = self.r.device_buffer(5)
        assert nbytes >= self.max_rows * 4
        self.lens_local = torch.as_tensor(_DevArray(ptr, self.max_rows), device=self.dev)
        self.lens_all = torch.zeros(world * self.max_rows, dtype=torch.int32, device=self.dev)
        self.lens_host = torch.zeros(world * self.max_rows, dtype=torch.int32).pin_memory()
        # device time stamps of the last step: start, trace done, quantise done, lengths exchanged (step_times)
        self.marks = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        self.lens_ready = torch.cuda.Event()
        # the one host buffer all ranks copy their rows into
        self.cap = width * height * 26 + 2 * height + 64
        self.shm = None
        self._host = None
        self._open_host_buffer()
        # pixel gather (optional path)
        self.block_elems = self.max_rows * width * 4
        self.block = self.gathered = None
        self.stream_len = 0

    # ---- the shared host buffer ------------------------------------------------------------------------
    def _open_host_buffer(self):
        lib = self.pkg.api.lib()
        if self.world == 1:
            p = self.pkg.api._P()
            if lib.trt_alloc_host(self.cap, C.byref(p)):
                raise MemoryError("pinned host buffer")
            self._host = p.value
            return
        from multiprocessing import shared_memory
        name = [None]
        if self.rank == 0:
            self.shm = shared_memory.SharedMemory(create=True, size=self.cap)
            name[0] = self.shm.name
        self.dist.broadcast_object_list(name, src=0)
        if self.rank != 0:
            self.shm = shared_memory.SharedMemory(name=name[0])
            try:                       # the creator unlinks; attaching ranks must not (Python < 3.13 registers them)
                from multiprocessing import resource_tracker
                resource_tracker.unregister(self.shm._name, "shared_memory")
            except Exception:
                pass
        self._host = C.addressof(C.c_char.from_buffer(self.shm.buf))
        if lib.trt_host_register(self._host, self.cap):
            raise MemoryError("cudaHostRegister of the shared frame buffer failed")

    def host_ptr(self):
        return self._host

    def upload_scene(self, prims):
        self.r.upload_scene(prims)

    # ---- one frame ---------------------------------------------------------------------------------------
    def step(self, u, mode):
        """Trace this rank's rows, quantise them here, exchange the row lengths.  Asynchronous on the context's
        stream; `read_stream()` completes the frame on the host."""
        torch, dist = self.torch, self.dist
        with torch.cuda.stream(self.stream):
            self.marks[0].record(self.stream)
            self.r.trace_async(u)
            self.marks[1].record(self.stream)
            if mode != self.pkg.MODE_NONE:
                self.r.quantise_rows_async(mode)
            self.marks[2].record(self.stream)
            if mode != self.pkg.MODE_NONE:
                if self.world > 1:
                    dist.all_gather_into_tensor(self.lens_all, self.lens_local)
                    self.lens_host.copy_(self.lens_all, non_blocking=True)
                else:
                    self.lens_host.copy_(self.lens_local, non_blocking=True)
                self.lens_ready.record(self.stream)
            self.marks[3].record(self.stream)
        self.mode = mode

    def read_stream(self) -> int:
        """Every rank copies its rows to their place in the shared host buffer over its own link; returns the
        stream's length.  The frame is complete on the host once ALL ranks have returned (barrier by the caller)."""
        self.lens_ready.synchronize()
        lens = self.lens_host.numpy().reshape(self.world, self.max_rows)
        off = stream_row_offsets(lens, self.h, self.world, self.stripe)
        self.stream_len = int(off[-1])
        self.r.size = (self.w, self.h)
        self.r.read_stream_rows(self._host, self.cap, off, wait=True)
        return self.stream_len

    def stream_bytes(self) -> bytes:
        """the assembled frame stream (valid on every rank of the node after read_stream + a barrier)"""
        return C.string_at(self._host, self.stream_len)

    def gather_pixels(self):
        """The float frame on rank 0's GPU: NCCL gather of the packed float4 rows (block r = shard r's rows).  Returns
        (device tensor of world blocks, block stride in bytes) on rank 0 -- what trt_quantise_gathered reads."""
        torch, dist = self.torch, self.dist
        if self.block is None:
            self.block = torch.zeros(self.block_elems, dtype=torch.float32, device=self.dev)
            self.gathered = torch.zeros(self.world * self.block_elems, dtype=torch.float32, device=self.dev) if self.rank == 0 else None
        with torch.cuda.stream(self.stream):
            self.r.pack_shard_pixels(self.block.data_ptr(), self.block.numel() * 4)
            if self.world > 1:
                glist = list(self.gathered.view(self.world, self.block_elems).unbind(0)) if self.rank == 0 else None
                dist.gather(self.block, glist, dst=0)
                return self.gathered, self.block_elems * 4
            return self.block, self.block_elems * 4

    def step_times(self):
        """(trace, quantise of the own rows, exchange of the row lengths incl. waiting for the slowest rank) of the
        last step in ms; synchronises."""
        self.marks[3].synchronize()
        m = self.marks
        return m[0].elapsed_time(m[1]), m[2].elapsed_time(m[3]), m[1].elapsed_time(m[2])

    def close(self):
        lib = self.pkg.api.lib()
        # the tensor over the context's row-length buffer must go before the context frees that buffer
        self.torch.cuda.synchronize(self.dev)
        self.lens_local = self.lens_all = self.lens_host = self.block = self.gathered = None
        if self._host is not None:
            if self.world == 1:
                lib.trt_free_host(self._host)
            else:
                lib.trt_host_unregister(self._host)
            self._host = None
        self.r.close()
        if self.shm is not None:
            try:
                self.shm.close()
                if self.rank == 0:
                    self.shm.unlink()
            except Exception:
                pass
            self.shm = None
