"""Shared test plumbing.  `-m "not gpu"` runs on a CPU-only box (oracle, host logic, ABI surface);
`-m gpu` runs the parity tests proper through the C ABI on a B200."""
import ctypes as C
import glob
import importlib
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


def load_rtp():
    return importlib.import_module("ray-tracing-project_b200")


@pytest.fixture(scope="session")
def rtp():
    return load_rtp()


@pytest.fixture(scope="session")
def po():
    import pyoracle
    pyoracle.lib()
    return pyoracle


def golden_cases():
    return sorted(os.path.basename(p)[5:-4] for p in glob.glob(os.path.join(GOLDEN, "case_*.npz")))


def load_case(name):
    return dict(np.load(os.path.join(GOLDEN, f"case_{name}.npz")))


_mesh_cache = {}


def load_golden_mesh(model):
    """Mesh arrays as the reference's loader left them (tests/golden/make_golden.py)."""
    import pyoracle
    key = str(model).replace(".obj", "")
    if key not in _mesh_cache:
        z = np.load(os.path.join(GOLDEN, f"mesh_{key}.npz"))
        m = pyoracle.Mesh(z["verts"], z["normals"], z["faces"], z["face_mat"], z["materials"], None, z["shape"])
        m.set_boxes(z["boxes"], z["box_off"], z["box_faces"])
        m.face_normal_crc = int(z["face_normal_crc"][0])
        _mesh_cache[key] = m
    return _mesh_cache[key]


def bits_equal(a, b):
    """Bit-exact float comparison that treats NaN == NaN."""
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    return (a.view(np.int32) == b.view(np.int32)) | (np.isnan(a) & np.isnan(b))


HOSTSIM_DEFS = os.environ.get("HOSTSIM_DEFS", "")  # e.g. -DRTP_BVH4: the same checks on a kernel-variant build
HOSTSIM = os.path.join(ROOT, "tests", "hostsim", "libhostsim%s.so" % HOSTSIM_DEFS.replace("-D", "_").replace(" ", "").lower())


@pytest.fixture(scope="session")
def hostsim():
    """tests/hostsim: the product's __host__ __device__ code compiled for the CPU (test-only)."""
    csrc = os.path.join(ROOT, "ray-tracing-project_b200", "csrc")
    srcs = [os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp")] + [os.path.join(csrc, f) for f in
                                                                      ("scene_host.cpp", "bvh_build.cpp", "ref_boxes.cpp")]
    deps = srcs + glob.glob(os.path.join(csrc, "*.h"))
    if not os.path.exists(HOSTSIM) or any(os.path.getmtime(d) > os.path.getmtime(HOSTSIM) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-DRTP_TRAV_STATS"] + HOSTSIM_DEFS.split() + [
                               "-I/usr/local/cuda/include", "-o", HOSTSIM] + srcs + ["-lpthread"])
    H = C.CDLL(HOSTSIM)
    H.hostsim_create.restype = C.c_char_p
    return H


class HostSimScene:
    def __init__(self, H, rtp, mesh, with_boxes=False):
        self.H, self.rtp = H, rtp
        kw = dict(boxes=mesh.boxes, box_off=mesh.box_off, box_faces=mesh.box_faces) if with_boxes else {}
        ms, self._keep = rtp.make_mesh_struct(mesh.verts, mesh.normals, mesh.faces, mesh.face_mat, mesh.materials,
                                              mesh.shape, None, **kw)
        self.h = C.c_void_p()
        err = H.hostsim_create(C.byref(ms), C.byref(self.h))
        assert not err, err
        info = (C.c_int * 6)()
        H.hostsim_info(self.h, info)
        self.info = dict(zip(("n_tris", "n_nodes", "n_boxes", "depth", "flagged", "degenerate"), list(info)))
        self.n_faces = len(mesh.faces)

    def boxes(self):
        nb = self.info["n_boxes"]
        bb, off, bf = np.zeros((nb, 6), np.float32), np.zeros(nb + 1, np.int32), np.zeros(self.n_faces, np.int32)
        self.H.hostsim_get_boxes(self.h, bb.ctypes.data_as(C.c_void_p), off.ctypes.data_as(C.c_void_p),
                                 bf.ctypes.data_as(C.c_void_p))
        return bb, off, bf

    def layout(self):
        """Device layout as flatten_scene made it: BVH nodes, triangles in leaf order, slot -> face id."""
        nn, nt = self.info["n_nodes"], self.info["n_tris"]
        nodes, tris, face_of = np.zeros((nn, 16), np.float32), np.zeros((nt, 16), np.float32), np.zeros(nt, np.int32)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        self.H.hostsim_get_layout(self.h, vp(nodes), vp(tris), vp(face_of))
        return nodes, tris, face_of

    def render(self, w, h, view, max_depth, point, dirl, pixels=None):
        import pyoracle
        l, keep = self.rtp.make_lights_struct(point, dirl)
        self.H.hostsim_set_lights(self.h, C.byref(l))
        cam = self.rtp.make_camera_struct(w, h, view)
        px = pyoracle.all_pixels(w, h) if pixels is None else np.ascontiguousarray(pixels, np.int32)
        n = len(px)
        rgb, hit = np.zeros((n, 3), np.float32), np.zeros(n, np.int32)
        t, d, cnt = np.zeros(n, np.float32), np.zeros((n, 3), np.float32), np.zeros(4, np.int64)
        bg = self.rtp.BACKGROUND
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        self.H.hostsim_render(self.h, C.byref(cam), vp(px), C.c_longlong(n), max_depth, vp(bg), vp(rgb), vp(hit),
                              vp(t), vp(d), vp(cnt))
        return dict(rgb=rgb, hit=hit, t=t, dir=d, counts=cnt)

    def rays(self, rays, anyhit):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out, t = np.zeros(len(rays), np.int32), np.zeros(len(rays), np.float32)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        self.H.hostsim_rays(self.h, vp(rays), C.c_longlong(len(rays)), int(anyhit), vp(out), vp(t))
        return out, t

    def close(self):
        self.H.hostsim_destroy(self.h)


def has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def soup(po, seed, n_tris=160):
    """Triangle soup built to produce ties: vertices snapped to a coarse grid (coplanar faces, shared edges and
    vertices that rays hit exactly), exact duplicates of faces, zero-area faces; several BoundingBoxes."""
    rng = np.random.default_rng(seed)
    grid = rng.integers(-4, 5, (n_tris // 2, 3)).astype(np.float32) / 4
    verts = np.concatenate([grid, rng.uniform(-1, 1, (n_tris // 2, 3)).astype(np.float32)])
    verts = np.concatenate([verts, np.ones((len(verts), 1), np.float32)], 1)
    faces = rng.integers(0, len(verts), (n_tris, 3)).astype(np.int32)
    faces[: n_tris // 3] = rng.integers(0, len(grid), (n_tris // 3, 3))          # grid-only faces
    faces[n_tris // 3: n_tris // 3 + 12] = faces[:12]                              # duplicates (equal t: first visited wins)
    faces[-6:, 2] = faces[-6:, 1]                                                   # degenerate
    # every face has a material: faces without one (-1) shade with whatever the previous hit left in the reference's
    # member variables - order-dependent there, the documented deviation of the product
    face_mat = rng.integers(0, 2, n_tris).astype(np.int32)
    mats = np.array([[0.1, 0.1, 0.1, 0.6, 0.3, 0.2, 0.9, 0.9, 0.9, 20, 1, 1],
                     [0.0, 0.1, 0.2, 0.2, 0.5, 0.7, 0.3, 0.3, 0.3, 3, 1, 1]], np.float32)
    m = po.Mesh(verts, None, faces, face_mat, mats)
    return m.build_boxes(min_faces=24)


SOUP_VIEWS = [None,                                                     # default camera: axis-aligned central rays
              np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, -1.5], np.float32),
              np.array([0, 0, 1, 0.25, 0, 1, 0, -0.25, -1, 0, 0, -2], np.float32)]
SOUP_LIGHTS = np.array([[-0.5, 2, 3, 1, 1, 1], [1.5, -1, 2, 0.5, 0.4, 0.3]], np.float32)
SOUP_DIRL = np.array([[0.2, -1.0, 0.1, 0.3, 0.3, 0.3]], np.float32)
