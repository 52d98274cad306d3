"""CPU: the host side of the multi-GPU tile partition, world_size 2 over gloo (no GPU needed): partition arithmetic,
the opaque-handle broadcast that carries the CUDA IPC handles, and the band gather of the NCCL fallback path."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from triangles_sdf_cpu_raytracing_b200 import multigpu as M


@pytest.mark.parametrize("H,band,count", [(1080, 32, 2), (1080, 32, 8), (2160, 32, 4), (90, 7, 3), (5, 32, 8), (64, 1, 2)])
def test_partition_covers_every_row_once(H, band, count):
    assert M.partition_is_exact(H, band, count)
    sizes = [M.band_rows_of(H, band, count, i).size for i in range(count)]
    assert sum(sizes) == H
    if H >= band * count * 4:
        assert max(sizes) - min(sizes) <= band  # interleaving balances the ranks


def test_single_rank_partition_is_the_whole_frame():
    assert np.array_equal(M.band_rows_of(17, 32, 1, 0), np.arange(17))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, H, W, band):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. the 64-byte handle exchange (what carries cudaIpcMemHandle_t from the root to every rank)
        payload = bytes(range(64)) if rank == 0 else None
        got = M.exchange_bytes(payload, 0)
        assert got == bytes(range(64))
        # 2. every rank "renders" its bands; the root assembles the frame exactly as the NCCL gather does
        rows = M.band_rows_of(H, band, world, rank)
        frame = np.zeros((H, W), np.float32)
        frame[rows] = (rows[:, None] * 1000 + np.arange(W)[None, :] + 0.5 * rank).astype(np.float32)
        block = torch.from_numpy(M.pack_bands(frame, rows))
        if rank == 0:
            for r in range(1, world):
                rr = M.band_rows_of(H, band, world, r)
                buf = torch.empty((rr.size, W), dtype=torch.float32)
                dist.recv(buf, src=r)
                M.unpack_bands(frame, rr, buf.numpy())
            want = np.zeros((H, W), np.float32)
            for r in range(world):
                rr = M.band_rows_of(H, band, world, r)
                want[rr] = (rr[:, None] * 1000 + np.arange(W)[None, :] + 0.5 * r).astype(np.float32)
            assert np.array_equal(frame, want)
        else:
            dist.send(block, dst=0)
        # 3. max-over-ranks timing reduction used by bench.py
        t = torch.tensor([float(rank + 1)])
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.item() == world
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_tile_gather():
    mp.spawn(_worker, args=(2, _free_port(), 45, 16, 8), nprocs=2, join=True)


def test_partition_matches_the_library_arithmetic_property():
    """Property test (hypothesis): for any frame height / band size / rank count, the rows a rank owns according to the
    host module are exactly the rows the library's launch arithmetic visits: make_frame sizes the launch to
    rows_here = my_bands * band_rows virtual rows (csrc/rt_api.cu) and the kernels map virtual row v to frame row
    (v / band_rows * band_count + band_index) * band_rows + v % band_rows, dropping rows >= H (csrc/rt_kernels.cuh)."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=300, deadline=None)
    @given(st.integers(1, 4400), st.integers(1, 64), st.integers(2, 8))
    def check(H, band, count):
        seen = np.zeros(H, np.int32)
        n_bands = (H + band - 1) // band
        for i in range(count):
            mine = (n_bands - i + count - 1) // count if n_bands > i else 0
            v = np.arange(mine * band)
            rows = (v // band * count + i) * band + v % band
            rows = rows[rows < H]
            assert np.array_equal(rows, M.band_rows_of(H, band, count, i))
            seen[rows] += 1
        assert np.all(seen == 1)

    check()
