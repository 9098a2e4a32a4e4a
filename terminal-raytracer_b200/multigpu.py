"""
Row-sharded multi-GPU frames: one process per GPU (torch.distributed), one trt context per process.

The path shards by pixels with no exchange while tracing (every pixel depends only on (x, y, uniforms,
scene): shader.wgsl:499-621).  Rows are dealt to ranks in stripes (trt_config.shard_*); RNG seeds and
pixel indices use global coordinates, so the union of the shards is bit-identical to a 1-GPU frame.
The only collective is the per-frame combine: an NCCL gather of each rank's packed pixel rows to the
root, which then quantises the assembled frame on its GPU (trt_quantise_gathered) -- north_star:
"NCCL reduce/gather only for that combine step".

`shard_rows` / `assemble_rows` are the host-side statement of the layout, used by the gloo tests.
"""
from __future__ import annotations

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


class ShardedRenderer:
    """One rank's part of a row-sharded frame.  Needs torch.distributed initialised (NCCL on GPUs)."""

    def __init__(self, pkg, width, height, rank, world, stripe_rows=4, device=0, conv=0):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.pkg = torch, dist, pkg
        self.w, self.h, self.rank, self.world, self.stripe = width, height, rank, world, stripe_rows
        self.r = pkg.Renderer(width, height, device=device, conv=conv, shard_index=rank, shard_count=world,
                              stripe_rows=stripe_rows)
        self.stream = torch.cuda.ExternalStream(self.r.cuda_stream(), device=device)
        self.max_rows = max_rows_per_shard(height, world, stripe_rows)
        self.block_elems = self.max_rows * width * 4
        dev = torch.device("cuda", device)
        self.block = torch.zeros(self.block_elems, dtype=torch.float32, device=dev)
        self.gathered = (torch.zeros(world * self.block_elems, dtype=torch.float32, device=dev) if rank == 0 else None)
        # device time stamps of the last step: start, trace done, gather done, quantise done (step_times)
        self.marks = [torch.cuda.Event(enable_timing=True) for _ in range(4)]

    def upload_scene(self, prims):
        self.r.upload_scene(prims)

    def step(self, u, mode):
        """Trace this rank's rows, gather to rank 0, quantise there.  Asynchronous on the context's stream."""
        torch, dist = self.torch, self.dist
        with torch.cuda.stream(self.stream):
            self.marks[0].record(self.stream)
            self.r.trace_async(u)
            self.marks[1].record(self.stream)
            self.r.pack_shard_pixels(self.block.data_ptr(), self.block.numel() * 4)
            if self.world > 1:
                glist = list(self.gathered.view(self.world, self.block_elems).unbind(0)) if self.rank == 0 else None
                dist.gather(self.block, glist, dst=0)
                src = self.gathered
            else:
                src = self.block
            self.marks[2].record(self.stream)
            if self.rank == 0 and mode != self.pkg.MODE_NONE:
                self.r.quantise_gathered(src.data_ptr(), self.block_elems * 4, self.w, self.h, self.world, self.stripe,
                                         mode, want_len=False)
            self.marks[3].record(self.stream)

    def step_times(self):
        """(trace, pack + gather incl. waiting for the slowest rank, quantise) of the last step in ms; synchronises."""
        self.marks[3].synchronize()
        m = self.marks
        return m[0].elapsed_time(m[1]), m[1].elapsed_time(m[2]), m[2].elapsed_time(m[3])

    def close(self):
        self.r.close()
