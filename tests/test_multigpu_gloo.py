"""World-size-2 test of the N > 1 path on CPU (gloo): row sharding, the per-frame gather of packed pixel
rows to rank 0 and the row reassembly that trt_quantise_gathered performs through its row map.

The CUDA kernels cannot run here, so each rank's rows are produced by the CPU oracle (test
infrastructure standing in for that rank's GPU); what is under test is the host-side plumbing of
terminal-raytracer_b200/multigpu.py: shard_rows / max_rows_per_shard / block layout / assemble_rows,
and that the union of shards is bit-identical to a single-rank frame (global pixel coordinates seed
the RNG, shader.wgsl:507)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, w, h, stripe, out_path):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import __graft_entry__ as g
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g.load_package()
    from terminal_raytracer_b200 import multigpu
    O = g.load_oracle()
    prims = O.flatten_scene(O.SCENE2)
    u = O.make_uniforms(O.SCENE2, len(prims), 77, 0, width=w, height=h, spp=8, max_depth=4)
    rows = multigpu.shard_rows(h, rank, world, stripe)
    idx = (rows[:, None].astype(np.uint32) * w + np.arange(w, dtype=np.uint32)[None, :]).reshape(-1)
    part = O.render(prims, u, idx=idx, threads=1)["pixels"]                 # this rank's rows, packed in local row order
    max_rows = multigpu.max_rows_per_shard(h, world, stripe)
    block = torch.zeros(max_rows * w * 4, dtype=torch.float32)
    block[: part.size] = torch.from_numpy(part.reshape(-1))
    gathered = [torch.zeros_like(block) for _ in range(world)] if rank == 0 else None
    dist.gather(block, gathered, dst=0)
    if rank == 0:
        frame = multigpu.assemble_rows([t.numpy() for t in gathered], w, h, world, stripe)
        np.save(out_path, frame)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("w,h,stripe", [(40, 23, 4), (33, 10, 1)])
def test_two_rank_gather_equals_single_rank_frame(O, tmp_path, w, h, stripe):
    import torch.multiprocessing as mp
    out = str(tmp_path / "frame.npy")
    port = 29000 + (os.getpid() + w) % 2000
    mp.spawn(_worker, args=(2, port, w, h, stripe, out), nprocs=2, join=True)
    frame = np.load(out)
    prims = O.flatten_scene(O.SCENE2)
    u = O.make_uniforms(O.SCENE2, len(prims), 77, 0, width=w, height=h, spp=8, max_depth=4)
    whole = O.render(prims, u)["pixels"]
    assert np.array_equal(frame.view(np.uint32), whole.view(np.uint32))
    # and therefore the root's byte stream equals the single-rank stream (lib.rs:503-532)
    assert O.quantise(frame, w, h, 1)[0] == O.quantise(whole, w, h, 1)[0]


def _stream_worker(rank, world, port, w, h, stripe, mode, out_path):
    """The row-addressed combine of multigpu.ShardedRenderer on CPU: every rank formats ITS rows (the oracle's
    quantiser standing in for the GPU's), the ranks all-gather the row lengths (gloo instead of NCCL) and every rank
    writes its rows at their offsets into one buffer (here: a file both ranks open)."""
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import __graft_entry__ as g
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g.load_package()
    from terminal_raytracer_b200 import multigpu
    O = g.load_oracle()
    prims = O.flatten_scene(O.SCENE2)
    u = O.make_uniforms(O.SCENE2, len(prims), 77, 0, width=w, height=h, spp=8, max_depth=4)
    rows = multigpu.shard_rows(h, rank, world, stripe)
    idx = (rows[:, None].astype(np.uint32) * w + np.arange(w, dtype=np.uint32)[None, :]).reshape(-1)
    part = O.render(prims, u, idx=idx, threads=1)["pixels"]
    stream, row_len = O.quantise(part, w, len(rows), mode)                  # this rank's rows back to back + their lengths
    max_rows = multigpu.max_rows_per_shard(h, world, stripe)
    mine = torch.zeros(max_rows, dtype=torch.int32)
    mine[: len(rows)] = torch.from_numpy(row_len.astype(np.int32))
    table = torch.zeros(world * max_rows, dtype=torch.int32)
    dist.all_gather_into_tensor(table, mine)
    lens = table.numpy().reshape(world, max_rows)
    off = multigpu.stream_row_offsets(lens, h, world, stripe)
    if rank == 0:
        with open(out_path, "wb") as f:
            f.truncate(int(off[-1]))
    dist.barrier()
    with open(out_path, "r+b") as f:                                         # "device -> host copies" of the own rows
        pos = 0
        for k, y in enumerate(rows):
            f.seek(int(off[y]))
            f.write(stream[pos: pos + int(row_len[k])])
            pos += int(row_len[k])
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("w,h,stripe,mode", [(40, 23, 4, 1), (33, 10, 1, 0), (17, 9, 2, 1)])
def test_two_rank_row_addressed_stream_equals_single_rank_stream(O, pkg, tmp_path, w, h, stripe, mode):
    import torch.multiprocessing as mp
    from terminal_raytracer_b200 import multigpu
    out = str(tmp_path / "stream.bin")
    port = 31000 + (os.getpid() + 7 * w + h) % 2000
    mp.spawn(_stream_worker, args=(2, port, w, h, stripe, mode, out), nprocs=2, join=True)
    prims = O.flatten_scene(O.SCENE2)
    u = O.make_uniforms(O.SCENE2, len(prims), 77, 0, width=w, height=h, spp=8, max_depth=4)
    whole, row_len = O.quantise(O.render(prims, u)["pixels"], w, h, mode)
    assert open(out, "rb").read() == whole
    # the pure-host statement of the same assembly
    lens, streams = [], []
    for r in range(2):
        rows = multigpu.shard_rows(h, r, 2, stripe)
        lens.append(row_len[rows])
        off = np.concatenate([[0], np.cumsum(row_len)]).astype(np.int64)
        streams.append(b"".join(whole[off[y]: off[y + 1]] for y in rows))
    assert multigpu.assemble_stream(streams, lens, h, 2, stripe) == whole
