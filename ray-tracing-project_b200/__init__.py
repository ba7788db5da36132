"""ray-tracing-project_b200: ctypes binding of the C ABI in include/rtp.h.

The product is the CUDA library `libflyscene_b200.so` (hand-written sm_100a kernels + the C++
host side).  This module only marshals numpy arrays into the ABI's plain structs for the tests
and bench.py; it contains no ray-tracing code and no CPU fallback.  If the library is missing or
no CUDA device is present, calls fail loudly.

Import with importlib (the directory name has hyphens):
    rtp = importlib.import_module("ray-tracing-project_b200")
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTP_LIB") or os.path.join(HERE, "libflyscene_b200.so")  # RTP_LIB: kernel-variant A/B runs

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)
_u8p = C.POINTER(C.c_uint8)

RAYS_SHADE, RAYS_CLOSEST, RAYS_ANYHIT = 0, 1, 2
DEFAULT_VIEW = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, -2], np.float32)  # Flycamera default
DEFAULT_LIGHT = np.array([[-0.5, 2.0, 3.0, 1, 1, 1]], np.float32)            # flyscene.cpp:37
BACKGROUND = np.array([0.9, 0.9, 0.9], np.float32)                           # flyscene.hpp:175
TILE = 32


class RtpMesh(C.Structure):
    _fields_ = [("n_verts", C.c_int32), ("n_faces", C.c_int32), ("n_mats", C.c_int32), ("n_boxes", C.c_int32),
                ("verts", _f32p), ("normals", _f32p), ("faces", _i32p), ("face_mat", _i32p),
                ("face_normal", _f32p), ("materials", _f32p), ("shape", C.c_float * 12),
                ("boxes", _f32p), ("box_off", _i32p), ("box_faces", _i32p),
                ("min_faces", C.c_int32), ("max_boxes", C.c_int32), ("build", C.c_int32), ("reserved", C.c_int32)]


class RtpCamera(C.Structure):
    _fields_ = [("view", C.c_float * 12), ("viewport", C.c_float * 4), ("fovy", C.c_float), ("aspect", C.c_float)]


class RtpLights(C.Structure):
    _fields_ = [("n_point", C.c_int32), ("n_dir", C.c_int32), ("point", _f32p), ("dir", _f32p)]


class RtpRenderOpts(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("max_depth", C.c_int32),
                ("background", C.c_float * 3), ("rank", C.c_int32), ("world", C.c_int32), ("flags", C.c_int32)]


class RtpStats(C.Structure):
    _fields_ = [("rays_primary", C.c_int64), ("rays_shadow", C.c_int64), ("rays_reflection", C.c_int64),
                ("hits", C.c_int64), ("ms_trace", C.c_float), ("ms_total", C.c_float),
                ("kernel_launches", C.c_int32), ("n_bvh_nodes", C.c_int32), ("n_triangles", C.c_int32),
                ("ms_closest", C.c_float), ("ms_anyhit", C.c_float), ("ms_shade", C.c_float),
                ("ms_paths", C.c_float), ("ms_camera", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


EXPORTS = ["rtp_scene_create", "rtp_scene_destroy", "rtp_set_lights", "rtp_set_camera", "rtp_render",
           "rtp_tile_buffer_bytes", "rtp_render_tiles_device", "rtp_untile_device", "rtp_trace_rays",
           "rtp_trace_rays_device", "rtp_scene_box_count", "rtp_scene_get_boxes", "rtp_scene_info",
           "rtp_last_error", "rtp_abi_version", "rtp_set_tile_groups", "rtp_get_tile_groups", "rtp_render_async", "rtp_render_wait", "rtp_peer_barrier", "rtp_peer_barrier_timeouts", "rtp_frame_stats", "rtp_tile_table", "rtp_render_image_device",
           "rtp_shared_alloc", "rtp_shared_open", "rtp_shared_free", "rtp_scene_create_multi", "rtp_scene_device_count",
           "rtp_device_count", "rtp_peer_barrier_reset", "rtp_trace_rays_device_noalloc"]

_lib = None


class RtpError(RuntimeError):
    pass


def lib():
    """Loads libflyscene_b200.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RtpError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(the CUDA library is the product; there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.rtp_last_error.restype = C.c_char_p
        L.rtp_tile_buffer_bytes.restype = C.c_size_t
        L.rtp_tile_buffer_bytes.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
        L.rtp_scene_create.argtypes = [C.POINTER(RtpMesh), C.c_int, C.POINTER(C.c_void_p)]
        L.rtp_scene_create_multi.argtypes = [C.POINTER(RtpMesh), C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]
        L.rtp_scene_device_count.argtypes = [C.c_void_p]
        L.rtp_device_count.argtypes = []
        L.rtp_peer_barrier_reset.argtypes = [C.c_void_p]
        L.rtp_scene_destroy.argtypes = [C.c_void_p]
        L.rtp_set_lights.argtypes = [C.c_void_p, C.POINTER(RtpLights)]
        L.rtp_set_camera.argtypes = [C.c_void_p, C.POINTER(RtpCamera)]
        L.rtp_set_tile_groups.argtypes = [C.c_void_p, C.c_int]
        L.rtp_get_tile_groups.argtypes = [C.c_void_p]
        L.rtp_render_async.argtypes = [C.c_void_p, C.POINTER(RtpRenderOpts), C.c_void_p, C.c_void_p]
        L.rtp_render_wait.argtypes = [C.c_void_p, C.POINTER(RtpStats)]
        L.rtp_peer_barrier.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p]
        L.rtp_peer_barrier_timeouts.argtypes = [C.c_void_p]
        L.rtp_render.argtypes = [C.c_void_p, C.POINTER(RtpRenderOpts), C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.POINTER(RtpStats)]
        L.rtp_render_tiles_device.argtypes = [C.c_void_p, C.POINTER(RtpRenderOpts), C.c_void_p, C.c_void_p,
                                              C.POINTER(RtpStats)]
        L.rtp_untile_device.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rtp_trace_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(RtpStats)]
        L.rtp_trace_rays_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.POINTER(RtpStats)]
        L.rtp_trace_rays_device_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p,
                                                    C.c_void_p, C.c_void_p]
        L.rtp_scene_box_count.argtypes = [C.c_void_p]
        L.rtp_scene_get_boxes.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rtp_scene_info.argtypes = [C.c_void_p, C.POINTER(RtpStats)]
        L.rtp_frame_stats.argtypes = [C.c_void_p, C.POINTER(RtpStats)]
        L.rtp_tile_table.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.rtp_render_image_device.argtypes = [C.c_void_p, C.POINTER(RtpRenderOpts), C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.POINTER(RtpStats)]
        L.rtp_shared_alloc.argtypes = [C.c_int, C.c_size_t, C.POINTER(C.c_void_p), C.c_void_p]
        L.rtp_shared_open.argtypes = [C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]
        L.rtp_shared_free.argtypes = [C.c_int, C.c_void_p, C.c_int]
        _lib = L
    return _lib


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a.reshape(shape) if shape else a


def _i32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a.reshape(shape) if shape else a


BUILD_HOST_SAH, BUILD_GPU_LBVH = 0, 1


def make_mesh_struct(verts, normals, faces, face_mat, materials, shape, face_normal=None, boxes=None,
                     box_off=None, box_faces=None, min_faces=0, max_boxes=0, build=0):
    """Builds an rtp_mesh over numpy arrays; returns (struct, keepalive list)."""
    verts, normals = _f32(verts, (-1, 4)), _f32(normals, (-1, 3))
    faces, face_mat = _i32(faces, (-1, 3)), _i32(face_mat, (-1,))
    materials = _f32(materials, (-1, 12))
    keep = [verts, normals, faces, face_mat, materials]
    m = RtpMesh()
    m.n_verts, m.n_faces, m.n_mats = len(verts), len(faces), len(materials)
    m.verts, m.normals = verts.ctypes.data_as(_f32p), normals.ctypes.data_as(_f32p)
    m.faces, m.face_mat = faces.ctypes.data_as(_i32p), face_mat.ctypes.data_as(_i32p)
    m.materials = materials.ctypes.data_as(_f32p)
    if face_normal is not None:
        face_normal = _f32(face_normal, (-1, 3))
        keep.append(face_normal)
        m.face_normal = face_normal.ctypes.data_as(_f32p)
    m.shape = (C.c_float * 12)(*_f32(shape, (12,)).tolist())
    if boxes is not None:
        boxes, box_off, box_faces = _f32(boxes, (-1, 6)), _i32(box_off), _i32(box_faces)
        keep += [boxes, box_off, box_faces]
        m.n_boxes = len(boxes)
        m.boxes, m.box_off, m.box_faces = (boxes.ctypes.data_as(_f32p), box_off.ctypes.data_as(_i32p),
                                           box_faces.ctypes.data_as(_i32p))
    m.min_faces, m.max_boxes, m.build = min_faces, max_boxes, build
    return m, keep


def make_camera_struct(w, h, view=None, fovy=60.0, aspect=None, viewport=None):
    c = RtpCamera()
    c.view = (C.c_float * 12)(*_f32(DEFAULT_VIEW if view is None else view, (12,)).tolist())
    vp = (0.0, 0.0, float(w), float(h)) if viewport is None else tuple(float(x) for x in viewport)
    c.viewport = (C.c_float * 4)(*vp)
    c.fovy = fovy
    c.aspect = float(np.float32(w) / np.float32(h)) if aspect is None else aspect  # flyscene.cpp:14
    return c


def make_lights_struct(point=None, dirl=None):
    point = _f32(DEFAULT_LIGHT if point is None else point, (-1, 6))
    dirl = _f32(np.zeros((0, 6)) if dirl is None else dirl, (-1, 6))
    l = RtpLights()
    l.n_point, l.n_dir = len(point), len(dirl)
    l.point, l.dir = point.ctypes.data_as(_f32p), dirl.ctypes.data_as(_f32p)
    return l, [point, dirl]


RENDER_PROFILE = 1


def make_opts(w, h, max_depth=2, background=BACKGROUND, rank=0, world=1, flags=0):
    o = RtpRenderOpts()
    o.width, o.height, o.max_depth, o.rank, o.world, o.flags = w, h, max_depth, rank, world, flags
    o.background = (C.c_float * 3)(*_f32(background, (3,)).tolist())
    return o


def device_count():
    n = lib().rtp_device_count()
    if n < 0:
        raise RtpError(f"rtp error {n}: {lib().rtp_last_error().decode(errors='replace')}")
    return n


class Scene:
    """A scene on one CUDA device (rtp_scene handle), or - devices=[...] - replicated on several devices of the box
    and driven from this one process (rtp_scene_create_multi: render() then shards the frame over them)."""

    def __init__(self, verts, normals, faces, face_mat, materials, shape, face_normal=None, boxes=None,
                 box_off=None, box_faces=None, device=-1, min_faces=0, max_boxes=0, build=0, devices=None):
        self._h = C.c_void_p()
        self._L = lib()
        m, keep = make_mesh_struct(verts, normals, faces, face_mat, materials, shape, face_normal, boxes,
                                   box_off, box_faces, min_faces, max_boxes, build)
        if devices is None:
            self._check(self._L.rtp_scene_create(C.byref(m), device, C.byref(self._h)))
        else:
            arr = (C.c_int * len(devices))(*[int(d) for d in devices])
            self._check(self._L.rtp_scene_create_multi(C.byref(m), arr, len(devices), C.byref(self._h)))
        self.n_faces = m.n_faces
        del keep

    def _check(self, rc):
        if rc != 0:
            raise RtpError(f"rtp error {rc}: {self._L.rtp_last_error().decode(errors='replace')}")

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.rtp_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_lights(self, point=None, dirl=None):
        l, keep = make_lights_struct(point, dirl)
        self._check(self._L.rtp_set_lights(self._h, C.byref(l)))

    def set_camera(self, w, h, view=None, fovy=60.0, aspect=None, viewport=None):
        c = make_camera_struct(w, h, view, fovy, aspect, viewport)
        self._check(self._L.rtp_set_camera(self._h, C.byref(c)))

    def set_tile_groups(self, groups):
        """rtp_set_tile_groups: bands of the frame that run on their own streams (0 = automatic)."""
        self._check(self._L.rtp_set_tile_groups(self._h, int(groups)))

    def tile_groups(self):
        return int(self._L.rtp_get_tile_groups(self._h))

    def info(self):
        st = RtpStats()
        self._check(self._L.rtp_scene_info(self._h, C.byref(st)))
        return st.as_dict()

    def frame_stats(self):
        """Waits for the last enqueued frame; ray counters (+ per-kernel ms with RENDER_PROFILE)."""
        st = RtpStats()
        self._check(self._L.rtp_frame_stats(self._h, C.byref(st)))
        return st.as_dict()

    def boxes(self):
        n = self._L.rtp_scene_box_count(self._h)
        b = np.zeros((n, 6), np.float32)
        off = np.zeros(n + 1, np.int32)
        bf = np.zeros(max(self.n_faces * 2, 1), np.int32)
        self._check(self._L.rtp_scene_get_boxes(self._h, b.ctypes.data, off.ctypes.data, bf.ctypes.data))
        return b, off, bf[:off[-1]]

    def render(self, w, h, max_depth=2, background=BACKGROUND, want=("rgb", "hit"), rank=0, world=1, out=None,
               flags=0):
        """rtp_render into host arrays. Returns dict(rgb, rgb8, hit, stats)."""
        o = make_opts(w, h, max_depth, background, rank, world, flags)
        out = dict(out or {})
        if "rgb" in want and "rgb" not in out:
            out["rgb"] = np.zeros((h, w, 3), np.float32)
        if "rgb8" in want and "rgb8" not in out:
            out["rgb8"] = np.zeros((h, w, 3), np.uint8)
        if "hit" in want and "hit" not in out:
            out["hit"] = np.full((h, w), -1, np.int32)
        st = RtpStats()
        ptr = lambda k: out[k].ctypes.data if k in out and k in want else None
        self._check(self._L.rtp_render(self._h, C.byref(o), ptr("rgb"), ptr("rgb8"), ptr("hit"), C.byref(st)))
        out["stats"] = st.as_dict()
        return out

    def render_async(self, w, h, rgb8_ptr=None, hit_ptr=None, max_depth=2, background=BACKGROUND):
        """rtp_render_async: enqueue one frame into (page-locked) host buffers given by address; at most two outstanding."""
        o = make_opts(w, h, max_depth, background, 0, 1, 0)
        self._check(self._L.rtp_render_async(self._h, C.byref(o), rgb8_ptr, hit_ptr))

    def render_wait(self):
        """rtp_render_wait: blocks until the oldest outstanding frame is in its host buffers; returns its stats."""
        st = RtpStats()
        self._check(self._L.rtp_render_wait(self._h, C.byref(st)))
        return st.as_dict()

    def peer_barrier(self, d_counter, target, timeout_ms=10000, stream=None):
        """rtp_peer_barrier: arrive at and wait on a counter in (peer) device memory, enqueued on `stream`."""
        self._check(self._L.rtp_peer_barrier(self._h, d_counter, int(target) & 0xffffffff, int(timeout_ms), stream))

    def device_count(self):
        return int(self._L.rtp_scene_device_count(self._h))

    def peer_barrier_reset(self):
        self._check(self._L.rtp_peer_barrier_reset(self._h))

    def peer_barrier_timeouts(self):
        n = self._L.rtp_peer_barrier_timeouts(self._h)
        if n < 0:
            self._check(n)
        return n

    def render_tiles_device(self, w, h, d_tiles_ptr, max_depth=2, background=BACKGROUND, rank=0, world=1,
                            stream=None, want_stats=False, flags=0):
        o = make_opts(w, h, max_depth, background, rank, world, flags)
        st = RtpStats()
        self._check(self._L.rtp_render_tiles_device(self._h, C.byref(o), d_tiles_ptr, stream,
                                                    C.byref(st) if want_stats else None))
        return st.as_dict() if want_stats else None

    def render_image_device(self, w, h, d_rgb8=None, d_hit=None, max_depth=2, background=BACKGROUND, rank=0, world=1,
                            stream=None, want_stats=False, flags=0):
        """rtp_render_image_device: this rank's tiles resolved straight into the finished image planes
        (which may be peer memory on another GPU)."""
        o = make_opts(w, h, max_depth, background, rank, world, flags)
        st = RtpStats()
        self._check(self._L.rtp_render_image_device(self._h, C.byref(o), d_rgb8, d_hit, stream,
                                                    C.byref(st) if want_stats else None))
        return st.as_dict() if want_stats else None

    def untile_device(self, w, h, world, d_all_ptr, rank_stride_bytes, d_rgb=None, d_rgb8=None, d_hit=None,
                      stream=None):
        self._check(self._L.rtp_untile_device(self._h, w, h, world, d_all_ptr, rank_stride_bytes, d_rgb, d_rgb8,
                                              d_hit, stream))

    def trace_rays(self, rays, mode, max_depth=2, background=BACKGROUND):
        rays = _f32(rays, (-1, 6))
        n = len(rays)
        st = RtpStats()
        bg = _f32(background, (3,))
        if mode == RAYS_CLOSEST:
            face, t = np.zeros(n, np.int32), np.zeros(n, np.float32)
            self._check(self._L.rtp_trace_rays(self._h, rays.ctypes.data, n, mode, max_depth, bg.ctypes.data, None,
                                               face.ctypes.data, t.ctypes.data, None, C.byref(st)))
            return dict(face=face, t=t, stats=st.as_dict())
        if mode == RAYS_ANYHIT:
            occ = np.zeros(n, np.uint8)
            self._check(self._L.rtp_trace_rays(self._h, rays.ctypes.data, n, mode, max_depth, bg.ctypes.data, None,
                                               None, None, occ.ctypes.data, C.byref(st)))
            return dict(occluded=occ, stats=st.as_dict())
        rgb, face = np.zeros((n, 3), np.float32), np.zeros(n, np.int32)
        self._check(self._L.rtp_trace_rays(self._h, rays.ctypes.data, n, mode, max_depth, bg.ctypes.data,
                                           rgb.ctypes.data, face.ctypes.data, None, None, C.byref(st)))
        return dict(rgb=rgb, face=face, stats=st.as_dict())

    def trace_rays_device(self, d_rays_ptr, n, mode, d_face=None, d_t=None, d_occ=None, stream=None, want_stats=False):
        st = RtpStats()
        self._check(self._L.rtp_trace_rays_device(self._h, d_rays_ptr, n, mode, d_face, d_t, d_occ, stream,
                                                  C.byref(st) if want_stats else None))
        return st.as_dict() if want_stats else None

    def trace_rays_device_noalloc(self, d_rays_ptr, n, mode, d_face=None, d_t=None, d_occ=None, stream=None):
        """rtp_trace_rays_device_noalloc: the same launches with the library's static scratch, nothing allocated."""
        self._check(self._L.rtp_trace_rays_device_noalloc(self._h, d_rays_ptr, n, mode, d_face, d_t, d_occ, stream))


SHARED_HANDLE_BYTES = 64


def _rc(rc, what):
    if rc != 0:
        raise RtpError(f"{what}: rtp error {rc}: {lib().rtp_last_error().decode(errors='replace')}")


def shared_alloc(device, nbytes):
    """(device pointer, 64-byte CUDA IPC handle) of a zero-filled buffer other processes can map."""
    ptr, handle = C.c_void_p(), C.create_string_buffer(SHARED_HANDLE_BYTES)
    _rc(lib().rtp_shared_alloc(device, nbytes, C.byref(ptr), handle), "rtp_shared_alloc")
    return ptr.value, handle.raw


def shared_open(device, handle):
    ptr = C.c_void_p()
    _rc(lib().rtp_shared_open(device, C.create_string_buffer(handle, SHARED_HANDLE_BYTES), C.byref(ptr)), "rtp_shared_open")
    return ptr.value


def shared_free(device, ptr, opened):
    _rc(lib().rtp_shared_free(device, ptr, 1 if opened else 0), "rtp_shared_free")


def tile_buffer_bytes(w, h, rank=0, world=1):
    return lib().rtp_tile_buffer_bytes(w, h, rank, world)


def tile_table(w, h, rank=0, world=1):
    """(tile ids owned by `rank`, local index of every tile within its owner's buffer)."""
    tx, ty = (w + TILE - 1) // TILE, (h + TILE - 1) // TILE
    tiles = np.zeros(tx * ty, np.int32)
    local = np.zeros(tx * ty, np.int32)
    n = lib().rtp_tile_table(w, h, rank, world, tiles.ctypes.data, local.ctypes.data)
    if n < 0:
        raise RtpError(f"rtp error {n}: {lib().rtp_last_error().decode()}")
    return tiles[:n], local
