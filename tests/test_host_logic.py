"""CPU: host-side logic -- generators, band sharding, the N>1 gather over gloo,
and the no-fallback rule (without a GPU the product refuses to run)."""
import os
import socket

import numpy as np
import pytest
import torch

from simpleraytracer_b200 import api, sharding


def test_generators_are_deterministic_and_exact():
    v1, a1, cam, light = api.gen_mesh(9, 12345, 64, 36)
    v2, a2, _, _ = api.gen_mesh(9, 12345, 64, 36)
    assert v1.shape == (12345, 9) and np.array_equal(v1, v2) and np.array_equal(a1, a2)
    v3, _, _, _ = api.gen_mesh(10, 12345, 64, 36)
    assert not np.array_equal(v1, v3)
    assert np.isfinite(v1).all() and (a1 >= 0).all() and (a1 <= 1).all()
    f, r, u = (np.array(cam.forward[:]), np.array(cam.right[:]), np.array(cam.up[:]))
    for x in (f, r, u):
        assert abs(np.linalg.norm(x) - 1) < 1e-6
    assert abs(f @ r) < 1e-6 and abs(f @ u) < 1e-6 and abs(r @ u) < 1e-6
    assert abs(cam.half_w / cam.half_h - 64 / 36) < 1e-5
    for n in (0, 1, 63, 64, 2000):
        assert api.gen_mesh(1, n, 8, 8)[0].shape == (n, 9)
    v, a, _, _ = api.gen_cornell_box(512, 512)
    assert v.shape == (12, 9)


@pytest.mark.parametrize("height,world", [(1080, 1), (1080, 2), (1080, 4), (1080, 8), (2160, 8), (27, 4), (5, 8)])
def test_row_bands_cover_the_frame(height, world):
    bands = sharding.row_bands(height, world)
    assert len(bands) == world and bands[0][0] == 0 and bands[-1][1] == height
    for (a0, a1), (b0, b1) in zip(bands, bands[1:]):
        assert a1 == b0 and a0 <= a1
    assert all(y0 % 8 == 0 for y0, _ in bands)
    sizes = [b - a for a, b in bands]
    assert max(sizes) - min(sizes) <= 8 or height < world * 8


@pytest.mark.parametrize("height,world", [(1080, 2), (1080, 8), (2160, 4), (13, 3), (8, 2)])
def test_interleaved_rows_partition_the_frame(height, world):
    owned = [sharding.interleaved_rows(height, world, r) for r in range(world)]
    assert sorted(y for rows in owned for y in rows) == list(range(height))
    for r, rows in enumerate(owned):
        assert all((y // 8) % world == r for y in rows)
    if height >= world * 64:
        sizes = [len(rows) for rows in owned]
        assert max(sizes) - min(sizes) <= 8


def _reduce_worker(rank, world, port, height, out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        frame = torch.zeros(height, 4, 3)
        rows = sharding.interleaved_rows(height, world, rank)
        frame[rows] = torch.tensor(rows, dtype=torch.float32)[:, None, None] + 0.5
        sharding.reduce_interleaved(frame, rank)
        if rank == 0:
            torch.save(frame, out)
    finally:
        dist.destroy_process_group()


def test_reduce_interleaved_world_size_2_gloo(tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "frame.pt")
    mp.spawn(_reduce_worker, args=(2, _free_port(), 43, out), nprocs=2, join=True)
    frame = torch.load(out)
    want = (torch.arange(43, dtype=torch.float32) + 0.5)[:, None, None].expand(43, 4, 3)
    assert torch.equal(frame, want)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _gather_worker(rank, world, port, height, out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        W = 6
        bands = sharding.row_bands(height, world)
        frame = torch.full((height, W, 3), -1.0)
        y0, y1 = bands[rank]
        ys = torch.arange(y0, y1, dtype=torch.float32)[:, None, None]
        frame[y0:y1] = ys * 100 + torch.arange(W, dtype=torch.float32)[None, :, None] + torch.arange(3) * 0.25
        sharding.gather_bands(frame, bands, rank)
        if rank == 0:
            torch.save(frame, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("height", [32, 27])   # equal bands (all_gather) and ragged bands (send/recv)
def test_gather_bands_world_size_2_gloo(tmp_path, height):
    import torch.multiprocessing as mp
    out = str(tmp_path / "frame.pt")
    mp.spawn(_gather_worker, args=(2, _free_port(), height, out), nprocs=2, join=True)
    frame = torch.load(out)
    ys = torch.arange(height, dtype=torch.float32)[:, None, None]
    want = ys * 100 + torch.arange(6, dtype=torch.float32)[None, :, None] + torch.arange(3) * 0.25
    assert torch.equal(frame, want)


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful on a box without a GPU")
def test_product_refuses_to_run_without_gpu():
    with pytest.raises(api.MLError):
        api.Context()


def test_oracle_side_generator_equals_the_product_generator():
    """oracle/libsrtscenes.so (scenes.cpp compiled alone by the host compiler) feeds the CPU arm of bench.py;
    it must emit exactly the arrays srtGenMesh / srtGenCornellBox emit."""
    import oracle_lib as O
    for seed, n, w, h in ((1, 100_000, 1920, 1080), (9, 12345, 64, 36), (3, 0, 8, 8)):
        v1, a1, c1, l1 = api.gen_mesh(seed, n, w, h)
        v2, a2, c2, l2 = O.gen_mesh(seed, n, w, h)
        assert np.array_equal(v1, v2) and np.array_equal(a1, a2)
        assert bytes(c1) == bytes(c2) and bytes(l1) == bytes(l2)
    v1, a1, c1, l1 = api.gen_cornell_box(512, 512)
    v2, a2, c2, l2 = O.gen_cornell_box(512, 512)
    assert np.array_equal(v1, v2) and np.array_equal(a1, a2) and bytes(c1) == bytes(c2) and bytes(l1) == bytes(l2)


@pytest.mark.parametrize("height,stride,requested", [(1080, 8, 4), (1080, 8, 16), (1080, 2, 4), (2160, 4, 1), (2160, 8, 16),
                                                      (360, 8, 4), (64, 8, 4), (7, 1, 3), (1080, 0, 8), (8, 8, 16)])
def test_chunk_layout_of_a_sharded_frame(height, stride, requested):
    """srtChunkLayout (include/srt_render.h; host arithmetic only, no device): the rows of a frame cut into at most
    `requested` flag chunks of whole (8 * stride)-row groups -- the layout srtRenderDeviceSignal's producers and
    srtGatherFrameHost's consumer must agree on.  Every chunk but the last holds the same number of 8-row blocks of
    every rank, the chunks tile the frame, and no chunk is empty."""
    from simpleraytracer_b200 import api
    p = api.make_params(1920, height, spp=1, row_block_stride=stride, row_block_phase=0)
    rows, n = api.chunk_layout(p, requested)
    group = 8 * max(stride, 1)
    assert rows % group == 0 and rows > 0
    assert 1 <= n <= max(requested, 1)
    assert (n - 1) * rows < height <= n * rows                  # tiles the frame, last chunk not empty
    per_rank = rows // group                                    # row blocks of each rank in a full chunk
    for k in range(n):
        y0, y1 = k * rows, min(height, (k + 1) * rows)
        blocks = [b for b in range(y0 // 8, (y1 + 7) // 8)]
        owners = [b % max(stride, 1) for b in blocks]
        counts = [owners.count(r) for r in range(max(stride, 1))]
        if k < n - 1:
            assert counts == [per_rank] * max(stride, 1)
        else:
            assert max(counts) <= per_rank and sum(counts) == len(blocks)
    # asking for more chunks than there are groups gives one chunk per group
    rows1, n1 = api.chunk_layout(p, 10_000)
    assert rows1 == group and n1 == (height + group - 1) // group
