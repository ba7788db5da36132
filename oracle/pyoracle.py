"""TEST INFRASTRUCTURE ONLY: ctypes bindings for oracle/liboracle.so (the plain-C CPU
restatement of the reference hot path) and a runner for oracle/_ref/ref_harness (the
reference's own code built headless by oracle/Makefile).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package never does.
"""
import ctypes as C
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")
REF_BIN = os.path.join(HERE, "_ref", "ref_harness")
REF_MODELS = os.path.join(HERE, "_ref", "resources", "models")

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int)


class _Scene(C.Structure):
    _fields_ = [("n_verts", C.c_int), ("n_faces", C.c_int), ("n_mats", C.c_int), ("n_boxes", C.c_int),
                ("verts", _f32p), ("normals", _f32p), ("faces", _i32p), ("face_mat", _i32p),
                ("face_normal", _f32p), ("materials", _f32p), ("shape", C.c_float * 12),
                ("boxes", _f32p), ("box_off", _i32p), ("box_faces", _i32p)]


class _Camera(C.Structure):
    _fields_ = [("view", C.c_float * 12), ("viewport", C.c_float * 4), ("fovy", C.c_float), ("aspect", C.c_float)]


class _Lights(C.Structure):
    _fields_ = [("n_point", C.c_int), ("n_dir", C.c_int), ("point", _f32p), ("dir", _f32p)]


class _Counts(C.Structure):
    _fields_ = [("primary", C.c_longlong), ("shadow", C.c_longlong), ("reflection", C.c_longlong),
                ("hits", C.c_longlong)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            subprocess.check_call(["make", "-C", HERE, "oracle"], stdout=subprocess.DEVNULL)
        _lib = C.CDLL(LIB_PATH)
        _lib.oracle_build_boxes.restype = C.c_int
    return _lib


def _fp(a):
    return a.ctypes.data_as(_f32p)


def _ip(a):
    return a.ctypes.data_as(_i32p)


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a.reshape(shape) if shape else a


def _i32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a.reshape(shape) if shape else a


DEFAULT_VIEW = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, -2], np.float32)  # flycamera.hpp:76-86,166-191
DEFAULT_LIGHT = np.array([[-0.5, 2.0, 3.0, 1, 1, 1]], np.float32)            # flyscene.cpp:37
BACKGROUND = np.array([0.9, 0.9, 0.9], np.float32)                           # flyscene.hpp:175


class Mesh:
    """Mesh arrays in the layout the reference's loader leaves them in (Tucano::Mesh)."""

    def __init__(self, verts, normals, faces, face_mat, materials, face_normal=None, shape=None):
        self.verts = _f32(verts, (-1, 4))
        self.faces = _i32(faces, (-1, 3))
        self.face_mat = _i32(face_mat, (-1,))
        self.materials = _f32(materials, (-1, 12))
        L = lib()
        if normals is None:
            normals = np.zeros((len(self.verts), 3), np.float32)
            L.oracle_compute_normals(_fp(self.verts), len(self.verts), _ip(self.faces), len(self.faces), _fp(normals))
        self.normals = _f32(normals, (-1, 3))
        if face_normal is None:
            face_normal = np.zeros((len(self.faces), 3), np.float32)
            L.oracle_face_normals(_fp(self.verts), _ip(self.faces), len(self.faces), _fp(face_normal))
        self.face_normal = _f32(face_normal, (-1, 3))
        if shape is None:
            shape = np.zeros(12, np.float32)
            L.oracle_shape_matrix(_fp(self.verts), len(self.verts), _fp(shape))
        self.shape = _f32(shape, (12,))
        self.boxes = self.box_off = self.box_faces = None

    def build_boxes(self, min_faces=300, max_boxes=2**31 - 1):
        L = lib()
        b, o, f = _f32p(), _i32p(), _i32p()
        n = L.oracle_build_boxes(_fp(self.verts), _ip(self.faces), len(self.faces), min_faces, max_boxes,
                                 C.byref(b), C.byref(o), C.byref(f))
        self.boxes = np.ctypeslib.as_array(b, (n, 6)).copy()
        self.box_off = np.ctypeslib.as_array(o, (n + 1,)).copy()
        self.box_faces = np.ctypeslib.as_array(f, (max(len(self.faces), 1),)).copy()[:len(self.faces)]
        for p in (b, o, f):
            L.oracle_free(p)
        return self

    def set_boxes(self, boxes, box_off, box_faces):
        self.boxes, self.box_off, self.box_faces = _f32(boxes, (-1, 6)), _i32(box_off), _i32(box_faces)
        return self

    def c_scene(self):
        if self.boxes is None:
            self.build_boxes()
        s = _Scene()
        s.n_verts, s.n_faces, s.n_mats, s.n_boxes = len(self.verts), len(self.faces), len(self.materials), len(self.boxes)
        s.verts, s.normals, s.faces, s.face_mat = _fp(self.verts), _fp(self.normals), _ip(self.faces), _ip(self.face_mat)
        s.face_normal, s.materials = _fp(self.face_normal), _fp(self.materials)
        s.shape = (C.c_float * 12)(*self.shape.tolist())
        s.boxes, s.box_off, s.box_faces = _fp(self.boxes), _ip(self.box_off), _ip(self.box_faces)
        return s


def c_camera(w, h, view=None, fovy=60.0, aspect=None):
    cam = _Camera()
    v = DEFAULT_VIEW if view is None else _f32(view, (12,))
    cam.view = (C.c_float * 12)(*v.tolist())
    cam.viewport = (C.c_float * 4)(0.0, 0.0, float(w), float(h))
    cam.fovy = fovy
    cam.aspect = np.float32(w) / np.float32(h) if aspect is None else aspect  # flyscene.cpp:14
    return cam


def c_lights(point=None, dirl=None):
    point = _f32(DEFAULT_LIGHT if point is None else point, (-1, 6))
    dirl = _f32(np.zeros((0, 6)) if dirl is None else dirl, (-1, 6))
    l = _Lights()
    l.n_point, l.n_dir, l.point, l.dir = len(point), len(dirl), _fp(point), _fp(dirl)
    l._keep = (point, dirl)
    return l


def all_pixels(w, h, stride=(1, 1), offset=(0, 0)):
    jj, ii = np.meshgrid(np.arange(offset[1], h, stride[1]), np.arange(offset[0], w, stride[0]), indexing="ij")
    return np.stack([ii.ravel(), jj.ravel()], 1).astype(np.int32)


def render(mesh, w, h, view=None, fovy=60.0, point_lights=None, dir_lights=None, max_depth=2,
           pixels=None, background=BACKGROUND, threads=1):
    """Oracle render of a pixel list. Returns dict(rgb, hit, t, dir, counts, pixels)."""
    L = lib()
    s = mesh.c_scene()
    cam = c_camera(w, h, view, fovy)
    li = c_lights(point_lights, dir_lights)
    px = all_pixels(w, h) if pixels is None else _i32(pixels, (-1, 2))
    n = len(px)
    rgb = np.zeros((n, 3), np.float32)
    hit = np.zeros(n, np.int32)
    t = np.zeros(n, np.float32)
    d = np.zeros((n, 3), np.float32)
    cnt = _Counts()
    bg = _f32(background, (3,))
    L.oracle_render(C.byref(s), C.byref(cam), C.byref(li), _ip(px), C.c_longlong(n), max_depth, _fp(bg),
                    threads, _fp(rgb), _ip(hit), _fp(t), _fp(d), C.byref(cnt))
    return dict(rgb=rgb, hit=hit, t=t, dir=d, pixels=px,
                counts=dict(primary=cnt.primary, shadow=cnt.shadow, reflection=cnt.reflection, hits=cnt.hits))


def shadow_rays(mesh, rays, threads=1):
    rays = _f32(rays, (-1, 6))
    out = np.zeros(len(rays), np.int32)
    s = mesh.c_scene()
    lib().oracle_shadow_rays(C.byref(s), _fp(rays), C.c_longlong(len(rays)), threads, _ip(out))
    return out


def closest_rays(mesh, rays, threads=1):
    rays = _f32(rays, (-1, 6))
    face = np.zeros(len(rays), np.int32)
    t = np.zeros(len(rays), np.float32)
    s = mesh.c_scene()
    lib().oracle_closest_rays(C.byref(s), _fp(rays), C.c_longlong(len(rays)), threads, _ip(face), _fp(t))
    return face, t


def trace_rays(mesh, rays, point_lights=None, dir_lights=None, max_depth=2, background=BACKGROUND):
    rays = _f32(rays, (-1, 6))
    rgb = np.zeros((len(rays), 3), np.float32)
    s = mesh.c_scene()
    li = c_lights(point_lights, dir_lights)
    bg = _f32(background, (3,))
    lib().oracle_trace_rays(C.byref(s), C.byref(li), _fp(rays), C.c_longlong(len(rays)), max_depth, _fp(bg), _fp(rgb))
    return rgb


def camera_derive(w, h, view=None, fovy=60.0):
    cam = c_camera(w, h, view, fovy)
    eye = np.zeros(3, np.float32)
    inv = np.zeros(12, np.float32)
    lib().oracle_camera_derive(C.byref(cam), _fp(eye), _fp(inv))
    return eye, inv


def quantise(rgb):
    rgb = _f32(rgb)
    out = np.zeros(rgb.size, np.uint8)
    lib().oracle_quantise(_fp(rgb), C.c_longlong(rgb.size), out.ctypes.data_as(C.POINTER(C.c_ubyte)))
    return out.reshape(rgb.shape)


# ---- the reference's own code (oracle/_ref/ref_harness) ------------------------------------
def ref_available():
    return os.path.exists(REF_BIN)


def ref_model(name):
    return os.path.join(REF_MODELS, name)


def ref_run(obj, w, h, view=None, fovy=60.0, point_lights=None, dir_lights=None, max_depth=2,
            pixels=None, stride=None, procs=1, dump_scene=False, chain=True, mode="render", rays=None,
            keep_dir=None, repeat=1, soft=None, srand=None):
    """Run the reference binary. Returns dict of numpy arrays (+ 'mesh' when dump_scene)."""
    if not ref_available():
        raise RuntimeError("oracle/_ref/ref_harness not built (run `make -C oracle ref` where /root/reference exists)")
    tmp = keep_dir or tempfile.mkdtemp(prefix="rtp_ref_")
    pre = os.path.join(tmp, "o")
    lines = [f"obj {os.path.abspath(obj)}", f"out {pre}", f"size {w} {h}", f"fovy {fovy!r}", f"max_depth {max_depth}",
             f"procs {procs}", f"dump_scene {int(dump_scene)}", f"chain {int(chain)}", f"mode {mode}",
             f"repeat {repeat}"]
    if view is not None:
        lines.append("view " + " ".join(repr(float(x)) for x in _f32(view, (12,))))
    if point_lights is not None or dir_lights is not None:
        lines.append("nolights")
        for l in _f32(point_lights if point_lights is not None else np.zeros((0, 6)), (-1, 6)):
            lines.append("light " + " ".join(repr(float(x)) for x in l))
        for l in _f32(dir_lights if dir_lights is not None else np.zeros((0, 6)), (-1, 6)):
            lines.append("dirlight " + " ".join(repr(float(x)) for x in l))
    if srand is not None:
        lines.append(f"srand {int(srand)}")
    for l in (soft if soft is not None else []):  # x y z r g b radius n: the reference's sphericalLight + addLight "s" branch
        lines.append("soft " + " ".join(repr(float(x)) for x in np.asarray(l, np.float32)))
    if pixels is not None:
        _i32(pixels, (-1, 2)).tofile(pre + ".pxin.i32")
        lines.append(f"pixel_file {pre}.pxin.i32")
    elif stride is not None:
        lines.append("stride %d %d %d %d" % tuple(stride))
    if rays is not None:
        _f32(rays, (-1, 6)).tofile(pre + ".rays.f32")
        lines.append(f"ray_file {pre}.rays.f32")
    with open(pre + ".job", "w") as f:
        f.write("\n".join(lines) + "\n")
    r = subprocess.run([REF_BIN, pre + ".job"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        raise RuntimeError("ref_harness failed:\n" + r.stdout.decode(errors="replace")[-2000:])
    out = {}

    def ld(suffix, dt, shape=None):
        p = pre + suffix
        if os.path.exists(p):
            a = np.fromfile(p, dt)
            return a.reshape(shape) if shape else a
        return None

    if mode == "render":
        out["rgb"], out["dir"] = ld(".rgb.f32", np.float32, (-1, 3)), ld(".dir.f32", np.float32, (-1, 3))
        out["pixels"] = ld(".pixels.i32", np.int32, (-1, 2))
        tt = open(pre + ".time.txt").read().split()
        out["trace_seconds"], out["wall_seconds"] = float(tt[0]), float(tt[1])
        out["counts"] = dict(primary=int(tt[4]), shadow=int(tt[5]), reflection=int(tt[6]), hits=int(tt[7]))
        out["pass_seconds"] = [float(x) for x in tt[8:]]
    else:
        tt = open(pre + ".time.txt").read().split()
        out["trace_seconds"] = float(tt[0])
    out["hit"], out["t"] = ld(".hit.i32", np.int32), ld(".t.f32", np.float32)
    out["lights"] = ld(".lights.f32", np.float32, (-1, 6))  # Flyscene::lights as traced (after sphericalLight)
    if dump_scene:
        m = Mesh(ld(".verts.f32", np.float32), ld(".normals.f32", np.float32), ld(".faces.i32", np.int32),
                 ld(".face_mat.i32", np.int32), ld(".materials.f32", np.float32), ld(".face_normal.f32", np.float32),
                 ld(".shape.f32", np.float32))
        m.set_boxes(ld(".boxes.f32", np.float32), ld(".box_off.i32", np.int32), ld(".box_faces.i32", np.int32))
        out["mesh"] = m
        out["camera"] = ld(".camera.f32", np.float32)
    if keep_dir is None:
        for fn in os.listdir(tmp):
            os.unlink(os.path.join(tmp, fn))
        os.rmdir(tmp)
    return out
