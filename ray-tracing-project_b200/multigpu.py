"""Multi-GPU frame driver: one process per GPU (torch.distributed), scene replicated, the image
split into interleaved 32x32 screen tiles (owner = (tx + ty) % world), ONE gather of the per-rank
tile-major buffers to rank 0 per frame (NCCL over NVLink on GPUs; the same code runs over gloo on
CPU tensors in the tests), then the library's un-tile kernel on rank 0.

This is plumbing only: buffers, the collective, call order.  The pixels come from
rtp_render_tiles_device / rtp_untile_device (include/rtp.h); nothing is computed here.
Replaces the row-band std::thread fan-out + join of the reference (src/flyscene.cpp:266-289).
"""
import importlib

import torch
import torch.distributed as dist


class TileGather:
    """Per-rank tile buffer + rank 0's gathered buffer, and the per-frame gather."""

    def __init__(self, w, h, rank, world, device, rtp=None):
        rtp = rtp or importlib.import_module("ray-tracing-project_b200")
        self.w, self.h, self.rank, self.world, self.device = w, h, rank, world, device
        self.stride_bytes = rtp.tile_buffer_bytes(w, h, rank, world)  # padded to the largest per-rank share
        n = self.stride_bytes // 4
        self.tiles = torch.zeros(n, dtype=torch.float32, device=device)
        self.gathered = torch.zeros(world * n, dtype=torch.float32, device=device) if (world > 1 and rank == 0) else None

    def gather(self):
        """All ranks call this after rendering their tiles; returns rank 0's [world][stride] buffer
        (the rank's own buffer when world == 1, None on the other ranks)."""
        if self.world == 1:
            return self.tiles
        dist.gather(self.tiles, list(self.gathered.chunk(self.world)) if self.rank == 0 else None, dst=0)
        return self.gathered


def render_frame(scene, tg, max_depth, stream, d_rgb8=None, d_rgb=None, d_hit=None, flags=0):
    """One frame on this rank: render own tiles, gather to rank 0, un-tile there."""
    scene.render_tiles_device(tg.w, tg.h, tg.tiles.data_ptr(), max_depth=max_depth, rank=tg.rank, world=tg.world,
                              stream=stream, flags=flags)
    src = tg.gather()
    if tg.rank == 0:
        scene.untile_device(tg.w, tg.h, tg.world, src.data_ptr(), tg.stride_bytes, d_rgb=d_rgb, d_rgb8=d_rgb8, d_hit=d_hit,
                            stream=stream)
