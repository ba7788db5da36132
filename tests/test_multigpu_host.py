"""Host-side logic of the multi-GPU path on CPU: world-size-2 (and 3) gloo process groups run the
same TileGather plumbing bench.py uses with NCCL; the tile tables come from the C ABI
(rtp_tile_table, no device needed).  Every rank fills its tile-major buffer with a pixel-id
pattern; after the gather rank 0 must be able to rebuild the full frame, each pixel exactly once.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT

TILE = 32


def pix_index(x, y):
    return (((y >> 2) * 4 + (x >> 3)) << 5) + ((y & 3) << 3) + (x & 7)


def _worker(rank, world, port, w, h, out_path):
    sys.path.insert(0, ROOT)
    import importlib
    rtp = importlib.import_module("ray-tracing-project_b200")
    mg = importlib.import_module("ray-tracing-project_b200.multigpu")
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        tg = mg.TileGather(w, h, rank, world, torch.device("cpu"), rtp)
        tiles, local = rtp.tile_table(w, h, rank, world)
        tx = (w + TILE - 1) // TILE
        buf = tg.tiles.numpy().reshape(-1, TILE * TILE, 4)
        buf[:] = -1.0
        assert len(tiles) <= buf.shape[0]
        for k, t in enumerate(tiles):  # what rtp_render_tiles_device writes: (r, g, b, face id) per pixel slot
            assert local[t] == k
            x0, y0 = (t % tx) * TILE, (t // tx) * TILE
            ys, xs = np.mgrid[0:TILE, 0:TILE]
            gx, gy = x0 + xs, y0 + ys
            inside = (gx < w) & (gy < h)
            slot = pix_index(xs, ys)
            buf[k, slot[inside], 0] = (gy * w + gx)[inside]
            buf[k, slot[inside], 1] = rank
        src = tg.gather()
        if rank == 0:
            np.save(out_path, src.numpy().reshape(world, -1, TILE * TILE, 4))
        else:
            assert src is None
    finally:
        dist.destroy_process_group()


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,w,h", [(2, 200, 150), (3, 97, 65), (2, 1920, 1080)])
def test_gather_rebuilds_the_frame(world, w, h, tmp_path):
    sys.path.insert(0, ROOT)
    import importlib
    rtp = importlib.import_module("ray-tracing-project_b200")
    out = str(tmp_path / "gathered.npy")
    mp.spawn(_worker, args=(world, free_port(), w, h, out), nprocs=world, join=True)
    allb = np.load(out)
    tx, ty = (w + TILE - 1) // TILE, (h + TILE - 1) // TILE
    _, local = rtp.tile_table(w, h, 0, world)
    owner_of = np.full(tx * ty, -1)
    for r in range(world):
        owner_of[rtp.tile_table(w, h, r, world)[0]] = r
    assert (owner_of >= 0).all()
    seen = np.zeros(w * h, np.int32)
    for t in range(tx * ty):  # the un-tile kernel's addressing, on the host
        owner = owner_of[t]
        x0, y0 = (t % tx) * TILE, (t // tx) * TILE
        for y in range(min(TILE, h - y0)):
            xs = np.arange(min(TILE, w - x0))
            v = allb[owner, local[t], pix_index(xs, y)]
            assert (v[:, 1] == owner).all()
            assert (v[:, 0] == (y0 + y) * w + x0 + xs).all()
            seen[(y0 + y) * w + x0 + xs] += 1
    assert (seen == 1).all()


def test_tile_tables_partition_and_balance():
    sys.path.insert(0, ROOT)
    import importlib
    rtp = importlib.import_module("ray-tracing-project_b200")
    for (w, h) in [(1920, 1080), (3840, 2160), (33, 31), (1, 1)]:
        tx, ty = (w + TILE - 1) // TILE, (h + TILE - 1) // TILE
        for world in (1, 2, 4, 8):
            owned = []
            for r in range(world):
                tiles, local = rtp.tile_table(w, h, r, world)
                assert np.array_equal(local[tiles], np.arange(len(tiles)))
                assert rtp.tile_buffer_bytes(w, h, r, world) >= len(tiles) * TILE * TILE * 16
                owned.append(tiles)
            allt = np.concatenate(owned)
            assert np.array_equal(np.sort(allt), np.arange(tx * ty))
            sizes = [len(o) for o in owned]
            assert max(sizes) - min(sizes) <= max(tx, ty)
    with pytest.raises(rtp.RtpError):
        rtp.tile_table(64, 64, 2, 2)
