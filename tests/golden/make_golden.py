"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/ref_harness, i.e.
/root/reference/src/flyscene.cpp + BoundingBox.cpp compiled headless by oracle/Makefile).

Run where /root/reference exists:   make -C oracle ref && python tests/golden/make_golden.py

Each fixture holds (a) the mesh exactly as the reference's OBJ loader + createFaces +
normalizeModelMatrix leave it, (b) the BoundingBox list generateBoundingBoxes built, and
(c) what Flyscene::traceRay returned for every pixel of a small frame: float RGB, primary-hit
face id, primary t and the primary ray directions.  The reference has no tests or golden
vectors of its own, so these are the pinned known answers.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
import pyoracle as po  # noqa: E402


def rotview(yaw, pitch, t):
    cy, sy, cp, sp = np.cos(yaw), np.sin(yaw), np.cos(pitch), np.sin(pitch)
    ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    rx = np.array([[1, 0, 0], [0, cp, -sp], [0, sp, cp]])
    v = np.zeros((3, 4), np.float32)
    v[:, :3] = (rx @ ry).astype(np.float32)
    v[:, 3] = np.float32(t)
    return v.ravel()


L4 = np.array([[-0.5, 2, 3, .4, .4, .4], [2, 2, 2, .4, .4, .4], [-2, 1, 1, .4, .4, .4], [0, 3, -1, .4, .4, .4]],
              np.float32)
D1 = np.array([[0.3, 0.8, 0.5, .5, .5, .5]], np.float32)

# name, model, w, h, view, max_depth, point lights, dir lights
CASES = [
    ("plane_c1", "plane.obj", 64, 64, None, 2, None, None),                       # BASELINE config 1 (reduced)
    ("plane_diag", "plane.obj", 96, 96, None, 3, L4[:2], None),                   # rays through the shared edge
    ("cube_rot", "cube.obj", 96, 64, rotview(0.6, 0.4, [0.1, -0.05, -2.2]), 3, None, None),
    ("cube_dirlight", "cube.obj", 64, 64, rotview(-0.5, 0.3, [0, 0, -2.0]), 2, L4[:1], D1),
    ("testding_2l", "testding.obj", 80, 80, rotview(-0.3, 0.2, [0, 0, -1.8]), 2, L4[:2], None),
    ("glas_nan", "glas.obj", 64, 64, None, 2, None, None),                        # Ns 506.25 -> NaN pixels (Q8)
    ("dodge_c1", "dodgeColorTest.obj", 64, 64, None, 2, None, None),              # the reference's default scene
    ("dodge_c2", "dodgeColorTest.obj", 96, 54, rotview(2.1, -0.3, [0.05, 0.1, -1.5]), 4, L4, None),  # C2 reduced
    # tests/golden/quirks.obj + quirks.mtl are authored for this repo (not reference files): CRLF, #vn != #v,
    # unknown usemtl, a quad + a 2-ref f line re-cut into triples, v/t/n forms, a tab-indented MTL key
    ("quirks", os.path.join(HERE, "quirks.obj"), 48, 48, None, 3, None, None),
]


# Soft lights: name, model, w, h, view, max_depth, srand seed, soft lights (x y z r g b radius n).
# The light list is drawn by the reference's own Flyscene::sphericalLight with libc rand() after srand(seed), then the
# light itself is appended with rgb / (n + 1) as the "s" branch of its addLight does: n + 1 point lights per entry.
SOFT_CASES = [
    ("soft17_dodge", "dodgeColorTest.obj", 64, 48, rotview(2.1, -0.3, [0.05, 0.1, -1.5]), 2, 7,
     [[-0.5, 2.0, 3.0, 0.9, 0.8, 0.7, 0.35, 16]]),                                               # L = 17
    ("soft65_testding", "testding.obj", 48, 48, rotview(-0.3, 0.2, [0, 0, -1.8]), 2, 11,
     [[1.0, 1.5, 2.0, 1.0, 1.0, 1.0, 0.5, 40], [-1.5, 2.0, 1.0, 0.6, 0.7, 0.9, 0.25, 23]]),       # L = 41 + 24 = 65
]


def main_soft():
    for name, model, w, h, view, md, seed, soft in SOFT_CASES:
        r = po.ref_run(po.ref_model(model), w, h, view=view, max_depth=md, soft=soft, srand=seed, procs=os.cpu_count() or 1,
                       dump_scene=True)
        np.savez_compressed(os.path.join(HERE, "case_" + name + ".npz"), model=model, w=w, h=h, view=view, max_depth=md,
                            srand=seed, soft=np.asarray(soft, np.float32), point_lights=r["lights"],
                            dir_lights=np.zeros((0, 6), np.float32), rgb=r["rgb"], hit=r["hit"], t=r["t"], dir=r["dir"],
                            camera=r["camera"],
                            counts=np.array([r["counts"][k] for k in ("primary", "shadow", "reflection", "hits")], np.int64))
        print(name, "lights", len(r["lights"]), "hits", int((r["hit"] >= 0).sum()), "shadow rays", r["counts"]["shadow"])


def main():
    meshes = {}
    for name, model, w, h, view, md, pl, dl in CASES:
        r = po.ref_run(model if os.path.isabs(model) else po.ref_model(model), w, h, view=view, max_depth=md, point_lights=pl, dir_lights=dl,
                       dump_scene=True)
        m = r["mesh"]
        model = os.path.basename(model)
        if model not in meshes:
            meshes[model] = m
            np.savez_compressed(os.path.join(HERE, "mesh_" + model.replace(".obj", "") + ".npz"),
                                verts=m.verts, normals=m.normals, faces=m.faces, face_mat=m.face_mat,
                                materials=m.materials, shape=m.shape, boxes=m.boxes, box_off=m.box_off,
                                box_faces=m.box_faces,
                                face_normal_crc=np.array([np.bitwise_xor.reduce(m.face_normal.view(np.uint32).ravel())]))
        np.savez_compressed(os.path.join(HERE, "case_" + name + ".npz"), model=model, w=w, h=h,
                            view=po.DEFAULT_VIEW if view is None else view, max_depth=md,
                            point_lights=po.DEFAULT_LIGHT if pl is None and dl is None else (np.zeros((0, 6), np.float32) if pl is None else pl),
                            dir_lights=np.zeros((0, 6), np.float32) if dl is None else dl,
                            rgb=r["rgb"], hit=r["hit"], t=r["t"], dir=r["dir"], camera=r["camera"],
                            counts=np.array([r["counts"][k] for k in ("primary", "shadow", "reflection", "hits")], np.int64))
        print(name, "hits", int((r["hit"] >= 0).sum()), "nan px", int(np.isnan(r["rgb"]).any(1).sum()))


if __name__ == "__main__":
    if "--soft-only" not in sys.argv:
        main()
    main_soft()
