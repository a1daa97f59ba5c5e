"""Multi-GPU frame partition, host side (voxelengine_b200/partition.py): band ownership and the gather to rank 0,
run as two real processes over gloo on CPU tensors (the GPU path uses the same code over NCCL)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from voxelengine_b200.partition import BAND_ROWS, FramePartition


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _pattern(W, H):
    """A frame whose every byte identifies its pixel row and column (mod 251)."""
    y, x, c = np.meshgrid(np.arange(H), np.arange(W), np.arange(4), indexing="ij")
    return ((y * 131 + x * 7 + c) % 251).astype(np.uint8)


def _worker(rank, world_size, port, W, H, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    part = FramePartition(W, H, world_size, rank)
    full = torch.from_numpy(_pattern(W, H))
    frame = part.alloc_frame("cpu")
    view = frame[: H * W * 4].view(H, W, 4)
    view.fill_(255)  # "not rendered"
    rows = part.rows_of(rank)
    view[rows] = full[rows]  # this rank "renders" exactly its own bands
    for _ in range(2):  # twice: the staging buffers are reused
        part.gather(frame)
    if rank == 0:
        np.save(out_path, view.numpy())
    # the pipelined variant: render callback per row chunk, asynchronous gathers
    view.fill_(255)
    mine = set(rows)

    def render_rows(y0, h):
        ys = [y for y in range(y0, y0 + h) if y in mine]
        view[ys] = full[ys]
    part.render_and_gather(frame, render_rows, n_chunks=3)
    if rank == 0:
        np.save(out_path + ".pipelined.npy", view.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("W,H,world_size", [(64, 40, 2), (48, 50, 2), (32, 8, 2), (40, 72, 3)])
def test_band_gather_over_gloo(tmp_path, W, H, world_size):
    out = str(tmp_path / "frame.npy")
    mp.spawn(_worker, args=(world_size, _free_port(), W, H, out), nprocs=world_size, join=True)
    assert np.array_equal(np.load(out), _pattern(W, H))
    assert np.array_equal(np.load(out + ".pipelined.npy"), _pattern(W, H))


def test_band_ownership_covers_every_row_once():
    for W, H, n in [(16, 2160, 8), (16, 4320, 8), (16, 37, 4), (16, 8, 8), (16, 1080, 3)]:
        seen = np.zeros(H, np.int32)
        for r in range(n):
            p = FramePartition(W, H, n, r)
            rows = p.rows_of(r)
            assert all(0 <= y < H for y in rows)
            seen[rows] += 1
            assert len(list(p.bands_of(r))) <= p.per_rank
            assert p.render_opts() == (n, r)
        assert (seen == 1).all()
        assert FramePartition(W, H, 1, 0).render_opts() == (0, 0)
        for k in (1, 2, 4, 7):
            ch = p.row_chunks(k)
            assert ch[0][0] == 0 and sum(h for _, h in ch) == H and len(ch) <= max(k, 1)
            assert all(y0 % (BAND_ROWS * n) == 0 for y0, _ in ch)
        assert p.frame_bytes() == ((H + BAND_ROWS - 1) // BAND_ROWS) * BAND_ROWS * W * 4


def test_rank_outside_world_is_rejected():
    with pytest.raises(ValueError):
        FramePartition(8, 8, 2, 2)


def test_checkerboard_partitions_refuse_row_copies():
    """A checkerboard frame's bands interleave in pixel space (thread row ty writes pixel rows 2 ty .. 2 ty + 2), so the
    paths that move whole pixel rows must refuse it instead of assembling stale rows."""
    part = FramePartition(64, 64, 2, 0, checkerboard=True)
    with pytest.raises(ValueError):
        part.gather(part.alloc_frame("cpu"))
    with pytest.raises(ValueError):
        part.render_and_gather(part.alloc_frame("cpu"), lambda y0, h: None, 2)
    from voxelengine_b200.partition import HostFrame
    with pytest.raises(ValueError):
        HostFrame(part, 0)
    single = FramePartition(64, 64, 1, 0, checkerboard=True)   # one rank: nothing to move, nothing to refuse
    assert single.gather(single.alloc_frame("cpu")) is not None
