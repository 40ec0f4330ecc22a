"""Row-band sharding of a frame over the ranks of one node.

Pixels are independent, the scene is replicated, so the render path needs no
data-path collective; the only exchange is the final gather of the bands, which
are contiguous slabs of the dense HWC frame.  torch.distributed is plumbing
here: NCCL on the GPUs, gloo in the CPU tests.
"""
from __future__ import annotations


def row_bands(height: int, world_size: int, align: int = 8):
    """Splits rows [0,height) into world_size contiguous bands whose starts are
    multiples of `align` (the BVH kernel's block height), as even as possible.
    Bands may be empty when height < world_size * align."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    units = (height + align - 1) // align
    bands = []
    for r in range(world_size):
        u0 = (units * r) // world_size
        u1 = (units * (r + 1)) // world_size
        bands.append((min(u0 * align, height), min(u1 * align, height)))
    return bands


def interleaved_rows(height: int, world_size: int, rank: int, block: int = 8):
    """Rows rank `rank` owns under interleaved sharding (srt_render_params.row_block_stride =
    world_size, row_block_phase = rank): 8-row blocks b = y // 8 with b % world_size == rank.
    Interleaving balances load when the cost per row varies over the frame (sky vs. geometry)."""
    return [y for y in range(height) if (y // block) % world_size == rank]


def reduce_interleaved(frame, rank, dst=0, group=None):
    """Fallback exchange for interleaved sharding when peer (IPC) writes are unavailable:
    every rank's frame holds its own rows and exact zeros elsewhere, so a sum-reduce onto
    `dst` assembles the frame bit-exactly (x + 0 == x)."""
    import torch.distributed as dist

    dist.reduce(frame, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return frame


def gather_bands(frame, bands, rank, dst=0, group=None):
    """Gathers every rank's band of `frame` (a [H, W, C] torch tensor whose rows
    [y0,y1) hold this rank's pixels) into rank `dst`'s frame, in place.
    Equal bands use one all_gather_into_tensor; ragged bands use send/recv."""
    import torch.distributed as dist

    world = len(bands)
    if world == 1:
        return frame
    sizes = {y1 - y0 for y0, y1 in bands}
    y0, y1 = bands[rank]
    if len(sizes) == 1 and bands[0][0] == 0 and bands[-1][1] == frame.shape[0]:
        dist.all_gather_into_tensor(frame, frame[y0:y1].contiguous(), group=group)
        return frame
    if rank == dst:
        reqs = []
        for r, (a, b) in enumerate(bands):
            if r != dst and b > a:
                reqs.append(dist.irecv(frame[a:b], src=r, group=group))
        for q in reqs:
            q.wait()
    elif y1 > y0:
        dist.send(frame[y0:y1].contiguous(), dst=dst, group=group)
    return frame
