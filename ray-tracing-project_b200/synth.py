"""Synthetic workloads for the BASELINE.json configs (host-side numpy only; no tracing here).

C3: ~1 M-triangle terrain / UV sphere;  C4: a shipped mesh instanced on a 25 x 25 grid
(~10 M triangles);  C5: ambient-occlusion style incoherent ray batches.
All generators are deterministic (seed 1234) and return arrays in the layout of rtp_mesh
(include/rtp.h): verts V x 4, normals V x 3, faces F x 3, face_mat F, materials M x 12, shape 12.
`write_obj` emits the same mesh as OBJ + MTL text so that the reference's own loader can read it.
"""
import os

import numpy as np

SEED = 1234
# one material: ka .2, kd .8, ks .3, Ns 10, Ni 1, d 1
MATERIAL_C3 = np.array([[.2, .2, .2, .8, .8, .8, .3, .3, .3, 10.0, 1.0, 1.0]], np.float32)
LIGHTS_C2 = np.array([[-0.5, 2, 3, .4, .4, .4], [2, 2, 2, .4, .4, .4], [-2, 1, 1, .4, .4, .4],
                      [0, 3, -1, .4, .4, .4]], np.float32)
LIGHT_DEFAULT = np.array([[-0.5, 2.0, 3.0, 1, 1, 1]], np.float32)


def shape_matrix(verts):
    """Mesh::loadVertices normalisation + Model::normalizeModelMatrix (tucano mesh.hpp:611-628,
    model.hpp:169-173): scale(1 / max extent) * translate(-bbox centre), 3x4 row-major."""
    v = np.asarray(verts, np.float32)[:, :3]
    lo, hi = v.min(0), v.max(0)
    ext = np.float32(np.abs(hi - lo).max())
    s = np.float32(1.0 / np.float64(ext))
    c = ((hi + lo).astype(np.float64) / 2.0).astype(np.float32)
    m = np.zeros((3, 4), np.float32)
    for r in range(3):
        m[r, r] = s
        m[r, 3] = np.float32(s * -c[r])
    return m.ravel()


def smooth_normals(verts, faces):
    """Normalised sum of unit face normals per vertex (objimporter.hpp:81-106)."""
    v = np.asarray(verts, np.float32)[:, :3]
    f = np.asarray(faces, np.int64)

    def unit(a):
        n = np.sqrt((a * a).sum(1, keepdims=True, dtype=np.float32))
        return np.where(n > 0, a / np.where(n > 0, n, 1), a).astype(np.float32)
    e0, e1 = unit(v[f[:, 1]] - v[f[:, 0]]), unit(v[f[:, 2]] - v[f[:, 0]])
    n = unit(np.cross(e0, e1).astype(np.float32))
    out = np.zeros_like(v)
    for k in range(3):
        np.add.at(out, f[:, k], n)
    return unit(out)


def _value_noise(x, y, seed):
    xi, yi = np.floor(x).astype(np.int64), np.floor(y).astype(np.int64)
    fx, fy = x - xi, y - yi

    def h(a, b):
        u = (a * 73856093) ^ (b * 19349663) ^ (seed * 83492791)
        u = (u ^ (u >> 13)) * 1274126177
        return ((u ^ (u >> 16)) & 0xFFFF) / 65535.0
    sx, sy = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy)
    a, b, c, d = h(xi, yi), h(xi + 1, yi), h(xi, yi + 1), h(xi + 1, yi + 1)
    return (a + (b - a) * sx) * (1 - sy) + (c + (d - c) * sx) * sy


def terrain(cells=708, seed=SEED, height=0.18):
    """Height field over the unit square in the xz-plane (y up), 4-octave value-noise fBm;
    2 * cells^2 triangles (708 -> 1 002 528)."""
    n = cells + 1
    g = np.linspace(0.0, 1.0, n)
    x, z = np.meshgrid(g, g, indexing="xy")
    y = np.zeros_like(x)
    amp, freq = 1.0, 4.0
    for o in range(4):
        y += amp * _value_noise(x * freq, z * freq, seed + o)
        amp *= 0.5
        freq *= 2.0
    y = (y / 1.875 - 0.5) * height
    verts = np.stack([x.ravel(), y.ravel(), z.ravel(), np.ones(n * n)], 1).astype(np.float32)
    i, j = np.meshgrid(np.arange(cells), np.arange(cells), indexing="xy")
    a = (j * n + i).ravel()
    b, c, d = a + 1, a + n, a + n + 1
    faces = np.concatenate([np.stack([a, c, b], 1), np.stack([b, c, d], 1)], 1).reshape(-1, 3).astype(np.int32)
    return _finish(verts, faces)


def uv_sphere(n=708):
    """n x n quads of a UV sphere -> 2 n^2 triangles; the pole rows hold zero-area triangles,
    which the reference never hits (their Face::normal is the zero vector)."""
    th = np.linspace(0.0, np.pi, n + 1)
    ph = np.linspace(0.0, 2 * np.pi, n + 1)
    t, p = np.meshgrid(th, ph, indexing="ij")
    verts = np.stack([np.sin(t) * np.cos(p), np.cos(t), np.sin(t) * np.sin(p), np.ones_like(t)], -1)
    verts = verts.reshape(-1, 4).astype(np.float32)
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="xy")
    a = (j * (n + 1) + i).ravel()
    b, c, d = a + 1, a + n + 1, a + n + 2
    faces = np.concatenate([np.stack([a, b, c], 1), np.stack([b, d, c], 1)], 1).reshape(-1, 3).astype(np.int32)
    m = _finish(verts, faces)
    nv = verts[:, :3] / np.linalg.norm(verts[:, :3], axis=1, keepdims=True)
    m["normals"] = nv.astype(np.float32)
    return m


def _finish(verts, faces, materials=MATERIAL_C3, face_mat=None):
    return dict(verts=verts, normals=smooth_normals(verts, faces), faces=faces,
                face_mat=np.zeros(len(faces), np.int32) if face_mat is None else face_mat,
                materials=np.asarray(materials, np.float32), shape=shape_matrix(verts))


def instanced(mesh, nx=25, nz=25, pitch=1.15):
    """C4: `mesh` (object space) baked nx * nz times on a grid in the xz-plane."""
    v = np.asarray(mesh["verts"], np.float32)
    lo, hi = v[:, :3].min(0), v[:, :3].max(0)
    ext = float((hi - lo).max()) * pitch
    nv, nf = len(v), len(mesh["faces"])
    verts = np.tile(v, (nx * nz, 1))
    normals = np.tile(np.asarray(mesh["normals"], np.float32), (nx * nz, 1))
    faces = np.tile(np.asarray(mesh["faces"], np.int32), (nx * nz, 1))
    k = 0
    for iz in range(nz):
        for ix in range(nx):
            verts[k * nv:(k + 1) * nv, 0] += np.float32(ix * ext)
            verts[k * nv:(k + 1) * nv, 2] += np.float32(iz * ext)
            faces[k * nf:(k + 1) * nf] += k * nv
            k += 1
    return dict(verts=verts, normals=normals, faces=faces,
                face_mat=np.tile(np.asarray(mesh["face_mat"], np.int32), nx * nz),
                materials=np.asarray(mesh["materials"], np.float32), shape=shape_matrix(verts))


def look_at(eye, target, up=(0, 1, 0)):
    """View matrix (3x4 row-major float32) of a camera at `eye` looking at `target` (-z forward)."""
    eye, target, up = (np.asarray(a, np.float64) for a in (eye, target, up))
    f = target - eye
    f /= np.linalg.norm(f)
    s = np.cross(f, up)
    s /= np.linalg.norm(s)
    u = np.cross(s, f)
    r = np.stack([s, u, -f])
    v = np.zeros((3, 4), np.float32)
    v[:, :3] = r.astype(np.float32)
    v[:, 3] = (-r @ eye).astype(np.float32)
    return v.ravel()


def write_obj(mesh, path, mtl_name="synthetic"):
    """OBJ + MTL text of `mesh` (v, vn per vertex, f) for the reference's own loader."""
    base = os.path.splitext(path)[0]
    mats = np.asarray(mesh["materials"], np.float32).reshape(-1, 12)
    with open(base + ".mtl", "w") as f:
        for i, m in enumerate(mats):
            f.write(f"newmtl {mtl_name}{i}\nNs {m[9]:.6f}\nKa {m[0]:.6f} {m[1]:.6f} {m[2]:.6f}\n"
                    f"Kd {m[3]:.6f} {m[4]:.6f} {m[5]:.6f}\nKs {m[6]:.6f} {m[7]:.6f} {m[8]:.6f}\nNi {m[10]:.6f}\nd {m[11]:.6f}\nillum 2\n")
    v, n, fc, fm = mesh["verts"], mesh["normals"], np.asarray(mesh["faces"]) + 1, np.asarray(mesh["face_mat"])
    with open(path, "w") as f:
        f.write(f"mtllib {os.path.basename(base)}.mtl\n")
        np.savetxt(f, v[:, :3], fmt="v %.9g %.9g %.9g")
        np.savetxt(f, n, fmt="vn %.9g %.9g %.9g")
        # one `usemtl` per run of equal material ids, faces kept in their original order (the
        # reference's loader starts a new index group at every usemtl: objimporter.hpp:168-193)
        bounds = np.flatnonzero(np.diff(fm)) + 1
        for seg in np.split(np.arange(len(fm)), bounds):
            if fm[seg[0]] >= 0:
                f.write(f"usemtl {mtl_name}{int(fm[seg[0]])}\n")
            np.savetxt(f, fc[seg], fmt="f %d %d %d")
    return path


# ---- C5: ambient-occlusion rays from a counter-based generator ------------------------------------
# Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11), written once over an array
# namespace: numpy on the host (tests) and torch on the device (bench.py) run the same integer arithmetic, so ray k of
# pixel p is a pure function of (seed, p, k) - the same ray set for every GPU count and in test and bench alike.
_PHILOX_M0, _PHILOX_M1, _PHILOX_W0, _PHILOX_W1, _M32 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85, 0xFFFFFFFF


def _mulhilo(m, x):
    """(high, low) 32 bits of the 64-bit product m * x; x: int64 array with values < 2^32.  The product may wrap in
    int64 - its two's-complement bits are still the product's bits."""
    p = x * m
    return (p >> 32) & _M32, p & _M32


def philox4x32(c0, c1, c2, c3, k0, k1, rounds=10):
    """Philox4x32: counter (c0..c3) and key (k0, k1) as int64 arrays / scalars holding 32-bit values -> 4 int64 arrays."""
    with np.errstate(over="ignore"):
        for _ in range(rounds):
            hi0, lo0 = _mulhilo(_PHILOX_M0, c0)
            hi1, lo1 = _mulhilo(_PHILOX_M1, c2)
            c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
            k0, k1 = (k0 + _PHILOX_W0) & _M32, (k1 + _PHILOX_W1) & _M32
    return c0, c1, c2, c3


def ao_directions(pixel_ids, N, n_per_point=64, seed=SEED, xp=np):
    """n_per_point directions per pixel, uniform on the hemisphere around the unit normal N[pixel]: Marsaglia's
    disc-rejection construction (only +, *, sqrt: bit-identical in numpy and torch), four attempts per direction from
    two Philox blocks keyed by (seed, 0) with counter (pixel id, ray k, block, 0); a direction whose four attempts all
    miss the disc (0.2 %) is the normal itself.  xp: numpy, or torch with pixel_ids / N on the device.
    Returns float32 [len(pixel_ids), n_per_point, 3]."""
    is_np = xp is np
    i64 = np.int64 if is_np else xp.int64
    f32 = np.float32 if is_np else xp.float32  # python float constants keep float32 arrays float32 in both libraries
    pix = pixel_ids.astype(i64)[:, None] if is_np else pixel_ids.to(i64)[:, None]
    k = (np.arange(n_per_point, dtype=i64) if is_np else xp.arange(n_per_point, dtype=i64, device=pixel_ids.device))[None, :]
    zero = pix * 0 + k * 0
    cand = []
    for block in range(2):
        r = philox4x32(pix + zero, k + zero, zero + block, zero, int(seed) & _M32, 0)
        u = [((v >> 8).astype(f32) if is_np else (v >> 8).to(f32)) * (2.0 ** -23) - 1.0 for v in r]  # 24 bits -> [-1, 1)
        cand += [(u[0], u[1]), (u[2], u[3])]
    where = np.where if is_np else xp.where
    # correctly rounded float32 square root in both libraries: torch's CPU float32 sqrt is an approximation, its
    # float64 one (CPU and CUDA) is IEEE, and rounding a 53-bit root to 24 bits is safe (53 >= 2 * 24 + 2)
    sqrt = np.sqrt if is_np else (lambda a: xp.sqrt(a.double()).float())
    done = None
    d = [N[:, None, c] + zero.astype(f32) if is_np else N[:, None, c] + zero.to(f32) for c in range(3)]  # fallback: N
    for x1, x2 in reversed(cand):  # the FIRST accepted attempt wins: apply them last to first
        s2 = x1 * x1 + x2 * x2
        ok = (s2 < 1.0) & (s2 > 0.0)
        root = sqrt(where(ok, 1.0 - s2, s2 * 0.0))
        c = (2.0 * x1 * root, 2.0 * x2 * root, 1.0 - 2.0 * s2)
        d = [where(ok, c[i], d[i]) for i in range(3)]
    dot = d[0] * N[:, None, 0] + d[1] * N[:, None, 1] + d[2] * N[:, None, 2]
    flip = dot < 0.0
    d = [where(flip, -d[i], d[i]) for i in range(3)]
    return np.stack(d, -1).astype(np.float32) if is_np else xp.stack(d, -1)


def ao_rays_pixels(P, N, pixel_ids, n_per_point=64, seed=SEED, offset=0.003, xp=np):
    """C5 as SURVEY.md 8(d) defines it: from every primary hit point P (moved off the surface by offset * N, N the
    interpolated normal) n_per_point any-hit rays towards ao_directions(pixel id): [len(P) * n_per_point, 6] float32
    (origin, direction) for rtp_trace_rays(ANYHIT) = Flyscene::shadow(origin, direction)."""
    d = ao_directions(pixel_ids, N, n_per_point, seed, xp)
    off = np.float32(offset) if xp is np else float(np.float32(offset))
    o = (P + off * N)[:, None, :] + d * 0.0
    if xp is np:
        return np.concatenate([o, d], 2).reshape(-1, 6).astype(np.float32)
    return xp.cat([o, d], 2).reshape(-1, 6).contiguous()


def world_vertices(mesh):
    """Mesh vertices through the shape (model) matrix, float32 [V, 3]."""
    shp = np.asarray(mesh["shape"], np.float32).reshape(3, 4)
    v = np.asarray(mesh["verts"], np.float32)[:, :3]
    return (v @ shp[:, :3].T + shp[:, 3]).astype(np.float32)


def shading_normals(mesh, face_ids, P):
    """Unit normal at P inside face face_ids: the vertex normals weighted by the sub-triangle areas opposite to them
    (what Flyscene::interpolateNormal computes, flyscene.cpp:572-600), without its inside test - P comes from a hit.
    Workload generation only (float32 numpy); falls back to the face normal where the weights vanish."""
    V = world_vertices(mesh)
    f = np.asarray(mesh["faces"])[face_ids]
    v0, v1, v2 = V[f[:, 0]], V[f[:, 1]], V[f[:, 2]]
    nv = np.asarray(mesh["normals"], np.float32)
    nv = nv / np.maximum(np.linalg.norm(nv, axis=1, keepdims=True), 1e-30)
    area = lambda a, b, c: np.linalg.norm(np.cross(b - a, c - a), axis=1, keepdims=True).astype(np.float32)
    P = np.asarray(P, np.float32)
    w0, w1, w2 = area(P, v1, v2), area(P, v2, v0), area(P, v0, v1)
    n = w0 * nv[f[:, 0]] + w1 * nv[f[:, 1]] + w2 * nv[f[:, 2]]
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    fn = np.cross(v1 - v0, v2 - v0)
    fn = fn / np.maximum(np.linalg.norm(fn, axis=1, keepdims=True), 1e-30)
    return np.where(ln > 1e-20, n / np.maximum(ln, 1e-30), fn).astype(np.float32)


def primary_hit_points(mesh, face_ids, eye, dirs):
    """Hit points of rays (eye, dirs) with the planes of faces face_ids (float32 numpy; workload generation)."""
    V = world_vertices(mesh)
    f = np.asarray(mesh["faces"])[face_ids]
    v0 = V[f[:, 0]]
    n = np.cross(V[f[:, 1]] - v0, V[f[:, 2]] - v0).astype(np.float32)
    eye = np.asarray(eye, np.float32)
    t = ((v0 - eye[None, :]) * n).sum(1) / (dirs * n).sum(1)
    return (eye[None, :] + t[:, None].astype(np.float32) * dirs).astype(np.float32)


def camera_rays(w, h, view, fovy=60.0, aspect=None):
    """(eye [3], unit directions [h * w, 3]) of the pixel corners (i, j), as Camera::screenToWorld + traceRayThread form
    them (camera.hpp:155-173, flyscene.cpp:301-308) - float32 numpy, not bit-exact (workload generation only)."""
    view = np.asarray(view, np.float64).reshape(3, 4)
    R, t = view[:, :3], view[:, 3]
    Rinv = np.linalg.inv(R)
    eye = (Rinv @ -t)
    scale = np.tan(np.radians(fovy / 2.0))
    aspect = w / h if aspect is None else aspect
    i, j = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64), indexing="xy")
    x = (2.0 * i / w - 1.0) * aspect * scale
    y = (1.0 - 2.0 * j / h) * scale
    pc = np.stack([x.ravel(), y.ravel(), -np.ones(w * h)], 1)
    pw = (pc - t[None, :]) @ Rinv.T
    d = pw - eye[None, :]
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return eye.astype(np.float32), d.astype(np.float32)


def ao_rays(P, N, n_per_point=64, seed=SEED, offset=0.003):
    """Round-1 generator (numpy's own Philox stream, rays depend on the ORDER of the points): kept for the incoherent
    closest-hit checks; C5 itself uses ao_rays_pixels above."""
    P, N = np.asarray(P, np.float32), np.asarray(N, np.float32)
    m = len(P)
    rng = np.random.Generator(np.random.Philox(key=seed))
    d = rng.standard_normal((m, n_per_point, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=2, keepdims=True)
    flip = (d * N[:, None, :]).sum(2, keepdims=True) < 0
    d = np.where(flip, -d, d)
    o = np.broadcast_to((P + offset * N)[:, None, :], d.shape)
    return np.concatenate([o, d], 2).reshape(-1, 6).astype(np.float32)
