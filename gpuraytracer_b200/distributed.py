"""Row tiles across ranks (one process per GPU) and the gather to rank 0.

The image is cut into tiles of ``tile_rows`` rows dealt round-robin to the ranks
(rank r owns tiles r, r+N, r+2N, ...): neighbouring rows cost about the same, so the
deal balances the load, where contiguous bands would not (the top of the Cornell box
hits the light and ends early, the borders miss the scene).  The kernel maps a
rank's compact local rows to global rows itself (RTLaunchParams.tileRows /
tileStride), so pixel, offset and camera indexing stay bit-identical to a one-GPU
render.  Rendering needs no exchange; the only collective is the final gather of
float4 rows to rank 0 (NCCL over NVLink on GPUs, gloo in the CPU tests).
"""
import numpy as np


def local_rows(height, rank, world, tile_rows=8):
    """Global row indices owned by `rank`, in local (compact) order."""
    rows = np.arange(height)
    return rows[(rows // tile_rows) % world == rank]


def tile_params(height, rank, world, tile_rows=8):
    """(row0, rows, tileRows, tileStride) for RTLaunchParams."""
    if world == 1:
        return 0, height, 0, 0
    n = int(local_rows(height, rank, world, tile_rows).size)
    return rank * tile_rows, n, tile_rows, tile_rows * world


def max_local_rows(height, world, tile_rows=8):
    return max(int(local_rows(height, r, world, tile_rows).size) for r in range(world))


def scatter_index(height, world, tile_rows=8):
    """index[y] = position of global row y inside the rank-major, padded gather buffer."""
    pad = max_local_rows(height, world, tile_rows)
    idx = np.empty(height, np.int64)
    for r in range(world):
        rows = local_rows(height, r, world, tile_rows)
        idx[rows] = r * pad + np.arange(rows.size)
    return idx, pad


_GATHER_CACHE = {}


def gather_rows(local, height, world, rank, tile_rows=8, group=None):
    """local: torch tensor [pad, W, C] (this rank's rows first, padding after).
    Returns the full [height, W, C] image on rank 0, None elsewhere.

    Rank 0 keeps, per image shape, the rank-major receive buffer, the row index on the device and the output image, so a
    step allocates nothing and moves every row twice (NCCL receive into the rank-major buffer, one index_select into
    place): the gather lists VIEWS of the receive buffer, there is no concatenation.  The returned image is reused by
    the next call of the same shape."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return local[:height]
    if rank != 0:
        dist.gather(local, None, dst=0, group=group)
        return None
    key = (tuple(local.shape), local.dtype, local.device, height, world, tile_rows)
    c = _GATHER_CACHE.get(key)
    if c is None:
        idx, pad = scatter_index(height, world, tile_rows)
        assert local.shape[0] == pad
        allrows = torch.empty((world * pad,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        c = dict(allrows=allrows, parts=list(allrows.view((world, pad) + tuple(local.shape[1:])).unbind(0)),
                 idx=torch.as_tensor(idx, device=local.device),
                 full=torch.empty((height,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device))
        _GATHER_CACHE.clear()                      # one shape at a time: these are frame-sized buffers
        _GATHER_CACHE[key] = c
    dist.gather(local, c["parts"], dst=0, group=group)
    torch.index_select(c["allrows"], 0, c["idx"], out=c["full"])
    return c["full"]
