"""A rank's 1/N share of the frame rendered ALONE on one GPU (virtual rank: rank 0 of world N): the floor the N-GPU
frame cannot go below, without any cross-GPU effect.  L2 flushed before every frame, CUDA events on the launch stream.
usage: python profiles/share_time.py [--workload c3|c2|c4] [--steps 30] [--worlds 1,2,4,8] [--timeline]"""
import argparse
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (workload definitions)

rtp = importlib.import_module("ray-tracing-project_b200")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--worlds", default="1,2,4,8")
    ap.add_argument("--all-ranks", action="store_true", help="time every rank's share (default: rank 0 and the last)")
    ap.add_argument("--timeline", action="store_true", help="with RTP_TIMELINE=1: print the launch timeline of the last frames (stderr)")
    a = ap.parse_args()
    wl = bench.build_workload(a.workload, importlib.import_module("ray-tracing-project_b200.synth"))
    m, w, h, md = wl["mesh"], wl["w"], wl["h"], wl["max_depth"]
    scene = rtp.Scene(m["verts"], m["normals"], m["faces"], m["face_mat"], m["materials"], m["shape"], device=0)
    scene.set_lights(wl["lights"], None)
    scene.set_camera(w, h, wl["view"])
    dev = torch.device("cuda:0")
    rgb8 = torch.zeros(w * h * 3, dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    out = {}
    for world in [int(x) for x in a.worlds.split(",")]:
        per_rank = []
        for rank in (range(world) if a.all_ranks else sorted({0, world - 1})):
            ms = []
            for it in range(a.steps + 3):
                flush.zero_()
                torch.cuda.synchronize()
                ev0.record()
                scene.render_image_device(w, h, d_rgb8=rgb8.data_ptr(), max_depth=md, rank=rank, world=world, stream=stream)
                ev1.record()
                torch.cuda.synchronize()
                if it >= 3:
                    ms.append(ev0.elapsed_time(ev1))
                if a.timeline and it >= a.steps + 1:
                    print(f"world {world} rank {rank} frame {it}: {ev0.elapsed_time(ev1):.4f} ms", file=sys.stderr, flush=True)
                    scene.frame_stats()  # RTP_TIMELINE=1: the library prints the end time of every launch of this frame
            st = scene.frame_stats()
            per_rank.append(dict(rank=rank, ms=float(np.mean(ms)), ms_min=float(np.min(ms)), launches=st["kernel_launches"],
                                 rays=st["rays_primary"] + st["rays_shadow"] + st["rays_reflection"], groups=scene.tile_groups()))
        out[world] = per_rank
        print(world, per_rank, flush=True)
    base = out[min(out)][0]["ms"]
    print(json.dumps(dict(workload=a.workload, shares=out,
                          efficiency_bound={k: base / (k * max(r["ms"] for r in v)) for k, v in out.items()})))


if __name__ == "__main__":
    main()
