"""Host-side multi-GPU logic on CPU: the interleaved row-tile deal, the kernel's
local->global row mapping (checked through the oracle, which implements the same
RTLaunchParams.tileRows/tileStride rule) and the gather to rank 0 (gloo, world 2 and 3)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_lib as O
from gpuraytracer_b200 import distributed as D


@pytest.mark.parametrize("height,world,tile", [(1080, 1, 8), (1080, 2, 8), (1080, 8, 8), (77, 3, 8), (2160, 4, 16), (5, 4, 8)])
def test_deal_is_a_partition(height, world, tile):
    seen = np.concatenate([D.local_rows(height, r, world, tile) for r in range(world)])
    assert sorted(seen.tolist()) == list(range(height))
    for r in range(world):
        row0, rows, tr, ts = D.tile_params(height, r, world, tile)
        mine = D.local_rows(height, r, world, tile)
        assert rows == mine.size
        if world > 1 and rows:
            ly = np.arange(rows)
            assert np.array_equal(row0 + (ly // tr) * ts + ly % tr, mine)     # the kernel's mapping
    idx, pad = D.scatter_index(height, world, tile)
    assert pad == D.max_local_rows(height, world, tile) and len(set(idx.tolist())) == height


def test_tiled_render_equals_full_frame():
    Wd, Ht, spp, world, tile = 64, 45, 3, 3, 4
    sc = O.cornell(Wd, Ht)
    full, _ = O.pathtrace(sc, Wd, Ht, spp)
    off_full = O.offsets(Wd, 0, Ht).reshape(Ht, Wd)
    idx, pad = D.scatter_index(Ht, world, tile)
    parts = []
    for r in range(world):
        row0, rows, tr, ts = D.tile_params(Ht, r, world, tile)
        mine = D.local_rows(Ht, r, world, tile)
        img, _ = O.pathtrace(sc, Wd, Ht, spp, row0=row0, rows=rows, tile_rows=tr, tile_stride=ts,
                             off=np.ascontiguousarray(off_full[mine]).reshape(-1))
        parts.append(np.concatenate([img, np.zeros((pad - rows, Wd, 4), np.float32)]))
    assert np.array_equal(np.concatenate(parts)[idx], full)


def _worker(rank, world, port, height, tile, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pad = D.max_local_rows(height, world, tile)
    mine = D.local_rows(height, rank, world, tile)
    local = torch.full((pad, 5, 4), -1.0)
    local[:mine.size] = torch.as_tensor(mine, dtype=torch.float32)[:, None, None]
    full = D.gather_rows(local, height, world, rank, tile)
    if rank == 0:
        q.put(bool(torch.equal(full[:, 0, 0], torch.arange(height, dtype=torch.float32))) and full.shape == (height, 5, 4))
    else:
        assert full is None
    dist.destroy_process_group()


@pytest.mark.parametrize("world,height", [(2, 1080), (3, 77)])
def test_gather_to_rank0_gloo(world, height):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, height, 8, q)) for r in range(world)]
    [p.start() for p in procs]
    ok = q.get(timeout=120)
    [p.join(60) for p in procs]
    assert ok and all(p.exitcode == 0 for p in procs)
