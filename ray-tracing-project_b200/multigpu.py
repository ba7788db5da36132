"""Multi-GPU frame driver: one process per GPU (torch.distributed), scene replicated, the image
split into interleaved 32x32 screen tiles (owner = (tx + ty) % world), ONE gather of the per-rank
tile-major buffers to rank 0 per frame (NCCL over NVLink on GPUs; the same code runs over gloo on
CPU tensors in the tests), then the library's un-tile kernel on rank 0.

Two ways to bring the tiles to GPU 0:
  * TileGather + render_frame: per-rank tile-major buffers, ONE NCCL gather per frame, un-tile kernel on
    rank 0 (also what the gloo CPU tests drive);
  * PeerImage + render_frame_peer: rank 0's image is mapped into every process through CUDA IPC and
    each rank's resolve kernel stores its finished pixels straight into it over NVLink
    (rtp_render_image_device) - the gather is fused into the last kernel of the frame; what is left
    of the collective is a barrier, and that is the library's own kernel too (rtp_peer_barrier).  The image is double-buffered so that rank 0 may still read
    frame f while the others already write frame f + 1.

This is plumbing only: buffers, the collective, call order.  The pixels come from the library
(include/rtp.h); nothing is computed here.
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


class PeerImage:
    """GPU 0's finished image (W x H x 3 uint8, and optionally W x H int32 face ids), double-buffered,
    mapped into every rank's address space through CUDA IPC (NVLink peer access)."""

    def __init__(self, w, h, rank, world, device_index, rtp=None, with_hit=False):
        self.rtp = rtp or importlib.import_module("ray-tracing-project_b200")
        self.w, self.h, self.rank, self.world, self.dev, self.with_hit = w, h, rank, world, device_index, with_hit
        self.rgb_bytes, self.hit_bytes = w * h * 3, (w * h * 4 if with_hit else 0)
        self.stride = (self.rgb_bytes + self.hit_bytes + 255) // 256 * 256
        handle = torch.zeros(self.rtp.SHARED_HANDLE_BYTES, dtype=torch.uint8, device=torch.device("cuda", device_index))
        self.opened = False
        if rank == 0:
            self.base, raw = self.rtp.shared_alloc(device_index, 2 * self.stride + 256)  # + the frame-barrier counter (zeroed)
            handle.copy_(torch.frombuffer(bytearray(raw), dtype=torch.uint8))
        if world > 1:
            dist.broadcast(handle, src=0)
        if rank != 0:
            self.base = self.rtp.shared_open(device_index, bytes(handle.cpu().numpy().tobytes()))
            self.opened = True
        self.frame = 0
        self.token = torch.zeros(1, dtype=torch.int32, device=torch.device("cuda", device_index))
        self.counter = self.base + 2 * self.stride  # uint32 in rank 0's memory, reached by every rank over NVLink
        self.start_counter = self.counter + 64      # a second counter: align_start
        self.starts = 0
        if world > 1:
            dist.barrier()  # every rank has the block mapped before anybody's first frame arrives at the counter

    def planes(self, frame=None):
        """(rgb8 pointer, hit pointer or None) of the buffer frame `frame` is written to."""
        f = self.frame if frame is None else frame
        p = self.base + (f & 1) * self.stride
        return p, (p + self.rgb_bytes if self.with_hit else None)

    def rgb8_tensor(self, frame):
        """The rgb8 plane of `frame` as a torch tensor over the shared memory (no copy)."""
        class _Span:
            pass
        sp = _Span()
        sp.__cuda_array_interface__ = dict(shape=(self.h, self.w, 3), typestr="|u1", data=(self.planes(frame)[0], False), version=2)
        return torch.as_tensor(sp, device=torch.device("cuda", self.dev))

    def close(self):
        if getattr(self, "base", None):
            self.rtp.shared_free(self.dev, self.base, self.opened)
            self.base = None


def align_start(scene, img, stream, timeout_ms=10000):
    """Device-side start line: enqueued on `stream` in front of a frame, every rank's GPU waits (rtp_peer_barrier on a
    second counter of the shared block) until all ranks' GPUs have reached it - what is enqueued behind it starts within
    a microsecond or two on all GPUs, whatever the skew between the ranks' host threads; the hosts are not blocked and
    keep enqueueing the frame meanwhile.  Measurement plumbing (bench.py places it before the start event of a timed
    step, after the host barrier): the timed region then begins at the same instant on every GPU."""
    if img.world > 1:
        img.starts += 1
        scene.peer_barrier(img.start_counter, img.world * img.starts, timeout_ms, stream)


def render_frame_peer(scene, img, max_depth, stream, flags=0, barrier="peer", timeout_ms=10000):
    """One frame: every rank renders its tiles and resolves them straight into rank 0's image (peer stores over
    NVLink), then arrives at the frame barrier.  barrier = "peer": the library's own one-thread kernel
    (rtp_peer_barrier: system-scope atomic + wait on a counter in rank 0's memory; no collective left on this route);
    "nccl": a 4-byte all-reduce as a stream-ordered barrier.  Either way work enqueued behind it on rank 0 sees
    everybody's pixels, no rank overwrites a buffer rank 0 may still read (the image is double-buffered and a rank
    cannot start frame f + 2 before every rank has finished frame f + 1), and the host is not blocked.
    Returns the (rgb8, hit) device pointers of the finished frame (meaningful on rank 0)."""
    d_rgb8, d_hit = img.planes()
    scene.render_image_device(img.w, img.h, d_rgb8=d_rgb8, d_hit=d_hit, max_depth=max_depth, rank=img.rank,
                              world=img.world, stream=stream, flags=flags)
    img.frame += 1
    if img.world > 1:
        if barrier == "peer":
            scene.peer_barrier(img.counter, img.world * img.frame, timeout_ms, stream)
        else:
            dist.all_reduce(img.token)
    return d_rgb8, d_hit
