"""Workload for the compute-sanitizer passes: C1 and C2-reduced through every launch sequence of
the library - synchronous frame, async frames, tile groups, ray batches (allocating and no-alloc entry points), the
multi-device scene on one GPU - without torch in the process.  Which kernels run is steered from outside by the
library's environment knobs (RTP_BUDGET / RTP_PARK_* force k_coop, RTP_DEEP picks k_paths or the per-level kernels).
usage: compute-sanitizer --tool memcheck|racecheck|synccheck|initcheck python profiles/sanitize_case.py"""
import ctypes as C
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
rtp = importlib.import_module("ray-tracing-project_b200")
from conftest import load_case, load_golden_mesh  # noqa: E402

rt = C.CDLL("/usr/local/cuda/lib64/libcudart.so")
rt.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
rt.cudaFree.argtypes = [C.c_void_p]


def dmalloc(n):
    p = C.c_void_p()
    assert rt.cudaMalloc(C.byref(p), n) == 0
    return p.value


def run(case, w, h, depth):
    c = load_case(case)
    m = load_golden_mesh(str(c["model"]))
    s = rtp.Scene(m.verts, m.normals, m.faces, m.face_mat, m.materials, m.shape, device=0)
    s.set_lights(c["point_lights"], c["dir_lights"] if len(c["dir_lights"]) else None)
    s.set_camera(w, h, c["view"])
    ref = None
    for groups in (1, 3):
        s.set_tile_groups(groups)
        out = s.render(w, h, max_depth=depth, want=("rgb8", "hit"))
        if ref is None:
            ref = out
        assert np.array_equal(out["hit"], ref["hit"]) and np.array_equal(out["rgb8"], ref["rgb8"])
    s.set_tile_groups(0)
    # async frames, two in flight
    bufs = [np.zeros((h, w, 3), np.uint8) for _ in range(2)]
    for k in range(4):
        if k >= 2:
            s.render_wait()
        s.render_async(w, h, bufs[k & 1].ctypes.data, None, max_depth=depth)
    s.render_wait(); s.render_wait()
    assert np.array_equal(bufs[0], ref["rgb8"].reshape(h, w, 3)) and np.array_equal(bufs[1], bufs[0])
    # ray batches: host API (all three modes), then device API with and without allocation
    rng = np.random.default_rng(5)
    n = 6000
    o = rng.uniform(-1.2, 1.2, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    rays = np.ascontiguousarray(np.concatenate([o, d], 1))
    a = s.trace_rays(rays, rtp.RAYS_CLOSEST)
    b = s.trace_rays(rays, rtp.RAYS_ANYHIT)
    s.trace_rays(rays[:500], rtp.RAYS_SHADE, max_depth=depth)
    d_rays, d_face, d_t, d_occ = dmalloc(rays.nbytes), dmalloc(4 * n), dmalloc(4 * n), dmalloc(n)
    assert rt.cudaMemcpy(d_rays, rays.ctypes.data, rays.nbytes, 1) == 0
    for _ in range(2):
        s.trace_rays_device_noalloc(d_rays, n, rtp.RAYS_CLOSEST, d_face=d_face, d_t=d_t)
        s.trace_rays_device_noalloc(d_rays, n, rtp.RAYS_ANYHIT, d_occ=d_occ)
    face, occ = np.zeros(n, np.int32), np.zeros(n, np.uint8)
    assert rt.cudaMemcpy(face.ctypes.data, d_face, 4 * n, 2) == 0 and rt.cudaMemcpy(occ.ctypes.data, d_occ, n, 2) == 0
    assert np.array_equal(face, a["face"]) and np.array_equal(occ, b["occluded"])
    for p in (d_rays, d_face, d_t, d_occ):
        rt.cudaFree(p)
    s.close()
    # the sharded frame in one process (two virtual devices on this GPU): peer stores + event join
    s2 = rtp.Scene(m.verts, m.normals, m.faces, m.face_mat, m.materials, m.shape, devices=[0, 0])
    s2.set_lights(c["point_lights"], c["dir_lights"] if len(c["dir_lights"]) else None)
    s2.set_camera(w, h, c["view"])
    out2 = s2.render(w, h, max_depth=depth, want=("rgb8", "hit"))
    assert np.array_equal(out2["hit"], ref["hit"]) and np.array_equal(out2["rgb8"], ref["rgb8"])
    s2.close()
    print(f"{case} {w}x{h} depth {depth}: ok")


run("plane_c1", 256, 256, 2)
run("dodge_c2", 320, 180, 4)
