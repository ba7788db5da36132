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


def ao_rays(P, N, n_per_point=64, seed=SEED, offset=0.003):
    """C5: for each hit point P with unit normal N, n_per_point directions uniform on the
    hemisphere around N from a counter-based generator (Philox, key = point index)."""
    P, N = np.asarray(P, np.float32), np.asarray(N, np.float32)
    m = len(P)
    rng = np.random.Generator(np.random.Philox(key=seed))
    d = rng.standard_normal((m, n_per_point, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=2, keepdims=True)
    flip = (d * N[:, None, :]).sum(2, keepdims=True) < 0
    d = np.where(flip, -d, d)
    o = np.broadcast_to((P + offset * N)[:, None, :], d.shape)
    return np.concatenate([o, d], 2).reshape(-1, 6).astype(np.float32)
