"""Host-side logic of the product (flatten, BoundingBox splitter, BVH build) and the shared
__host__ __device__ traversal / intersection / shading code, run on the CPU through
tests/hostsim (test-only build) and compared bit-for-bit with the reference goldens and the
oracle.  The CUDA kernels call exactly these functions; the GPU parity tests are in
test_gpu_parity.py."""
import numpy as np
import pytest

from conftest import HostSimScene, bits_equal, golden_cases, load_case, load_golden_mesh, soup, SOUP_VIEWS, SOUP_LIGHTS, SOUP_DIRL


@pytest.mark.parametrize("name", golden_cases())
def test_shared_device_code_matches_golden(hostsim, rtp, name):
    c = load_case(name)
    m = load_golden_mesh(c["model"])
    s = HostSimScene(hostsim, rtp, m)
    o = s.render(int(c["w"]), int(c["h"]), c["view"], int(c["max_depth"]), c["point_lights"], c["dir_lights"])
    s.close()
    assert bits_equal(o["dir"], c["dir"]).all()
    assert np.array_equal(o["hit"], c["hit"])
    assert bits_equal(o["t"], c["t"]).all()
    assert bits_equal(o["rgb"], c["rgb"]).all()
    assert o["counts"].tolist() == c["counts"].tolist()


@pytest.mark.parametrize("model", ["plane", "cube", "testding", "glas", "dodgeColorTest"])
def test_product_boundingbox_splitter_matches_reference(hostsim, rtp, model):
    m = load_golden_mesh(model)
    s = HostSimScene(hostsim, rtp, m, with_boxes=False)  # boxes built by csrc/ref_boxes.cpp
    bb, off, bf = s.boxes()
    s.close()
    assert bb.shape == m.boxes.shape and np.array_equal(bb.view(np.int32), m.boxes.view(np.int32))
    assert np.array_equal(off, m.box_off) and np.array_equal(bf, m.box_faces)


def test_caller_supplied_boxes_restrict_visibility(hostsim, rtp, po):
    """Faces missing from the BoundingBox list are invisible, as in flyscene.cpp:381-393."""
    m = load_golden_mesh("cube")
    keep = m.box_faces[::2].copy()
    m2 = po.Mesh(m.verts, m.normals, m.faces, m.face_mat, m.materials, None, m.shape)
    m2.set_boxes(m.boxes, np.array([0, len(keep)], np.int32), keep)
    s = HostSimScene(hostsim, rtp, m2, with_boxes=True)
    a = s.render(48, 48, None, 2, None, None)
    s.close()
    b = po.render(m2, 48, 48)
    assert np.array_equal(a["hit"], b["hit"]) and bits_equal(a["rgb"], b["rgb"]).all()
    assert set(np.unique(a["hit"])) <= set(keep.tolist()) | {-1}


def shrunken_boxes_mesh(po, model="dodgeColorTest", factor=0.8):
    """BoundingBoxes shrunk towards their centres: the reference's slab test (flyscene.cpp:484-507)
    now rejects boxes whose faces a ray does hit, so those faces become invisible (Q11/Q13) -
    exercises the per-candidate visibility filter and the deferred check + exact re-run."""
    m = load_golden_mesh(model)
    c = 0.5 * (m.boxes[:, :3] + m.boxes[:, 3:])
    b = np.concatenate([c + factor * (m.boxes[:, :3] - c), c + factor * (m.boxes[:, 3:] - c)], 1).astype(np.float32)
    m2 = po.Mesh(m.verts, m.normals, m.faces, m.face_mat, m.materials, None, m.shape)
    return m2.set_boxes(b, m.box_off, m.box_faces)


def test_reference_box_filter_is_reproduced(hostsim, rtp, po):
    m2 = shrunken_boxes_mesh(po)
    c = load_case("dodge_c2")
    s = HostSimScene(hostsim, rtp, m2, with_boxes=True)
    a = s.render(96, 54, c["view"], 3, c["point_lights"], None)
    s.close()
    b = po.render(m2, 96, 54, view=c["view"], max_depth=3, point_lights=c["point_lights"])
    full = po.render(load_golden_mesh("dodgeColorTest"), 96, 54, view=c["view"], max_depth=3, point_lights=c["point_lights"])
    assert (b["hit"] != full["hit"]).sum() > 50, "the shrunken boxes must actually change the image"
    assert np.array_equal(a["hit"], b["hit"]) and bits_equal(a["t"], b["t"]).all()
    assert bits_equal(a["rgb"], b["rgb"]).all()


def test_ray_batches_match_oracle(hostsim, rtp, po):
    m = load_golden_mesh("dodgeColorTest")
    rng = np.random.default_rng(7)
    n = 3000
    o = rng.uniform(-1.2, 1.2, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    d[:50, 0] = 0.0  # axis-parallel directions: inf / NaN slabs in the reference box test (Q13)
    d[50:100, 1] = 0.0
    d[100:120, :2] = 0.0
    rays = np.concatenate([o, d], 1)
    s = HostSimScene(hostsim, rtp, m)
    face, t = s.rays(rays, anyhit=False)
    occ, _ = s.rays(rays, anyhit=True)
    s.close()
    f2, t2 = po.closest_rays(m, rays, threads=4)
    assert np.array_equal(face, f2) and bits_equal(t, t2).all()
    assert np.array_equal(occ, po.shadow_rays(m, rays, threads=4))
    assert (face >= 0).sum() > 100


def test_tile_tables_partition_the_image(hostsim):
    import ctypes as C
    for (w, h, world) in [(64, 64, 1), (100, 70, 2), (1920, 1080, 8), (33, 31, 4)]:
        tx, ty = (w + 31) // 32, (h + 31) // 32
        seen = np.zeros(tx * ty, np.int32)
        for r in range(world):
            tiles, loc, n = np.zeros(tx * ty, np.int32), np.zeros(tx * ty, np.int32), C.c_int()
            hostsim.hostsim_tile_tables(w, h, r, world, tiles.ctypes.data_as(C.c_void_p), C.byref(n),
                                        loc.ctypes.data_as(C.c_void_p))
            own = tiles[:n.value]
            seen[own] += 1
            assert np.array_equal(loc[own], np.arange(n.value))
            assert abs(n.value - tx * ty / world) <= max(tx, ty)
        assert (seen == 1).all()


def test_scene_build_does_not_depend_on_the_thread_count(hostsim, rtp, po, monkeypatch):
    """The threaded host build (BoundingBox passes split concurrently, large boxes swept in chunks,
    per-face flatten and BVH subtrees as tasks) must produce, bit for bit, the arrays the
    single-threaded build produces - and the BoundingBox list of the oracle's sequential splitter.
    80 000 faces: the first passes take the chunked path (ref_boxes.cpp kChunkedSplitMin)."""
    import importlib
    from types import SimpleNamespace
    synth = importlib.import_module("ray-tracing-project_b200.synth")
    d = synth.terrain(200)
    # a NaN vertex and signed zeros: the min/max merge has to treat them like the sequential loop
    d["verts"][7, 1] = np.nan
    d["verts"][:3, 2] = np.float32(-0.0)
    m = SimpleNamespace(**d)
    built = []
    for threads in ("1", "8", "3"):
        monkeypatch.setenv("HOSTSIM_THREADS", threads)
        s = HostSimScene(hostsim, rtp, m)
        built.append((s.info, s.boxes(), s.layout()))
        s.close()
    ref = built[0]
    assert ref[0]["n_boxes"] > 300
    for info, boxes, layout in built[1:]:
        assert info == ref[0]
        for a, b in zip(boxes + layout, ref[1] + ref[2]):
            assert a.shape == b.shape and np.array_equal(a.view(np.int32), b.view(np.int32))
    pm = po.Mesh(d["verts"], d["normals"], d["faces"], d["face_mat"], d["materials"], None, d["shape"]).build_boxes()
    bb, off, bf = ref[1]
    assert np.array_equal(off, pm.box_off) and np.array_equal(bf, pm.box_faces) and bits_equal(bb, pm.boxes).all()


@pytest.mark.parametrize("seed", range(24))
def test_tie_heavy_random_scenes_match_the_oracle(hostsim, rtp, po, seed):
    """Fuzz of the hit / tie / visibility rules where no golden exists: the product's traversal
    (BVH order, lexicographic (t, rank) minimum, deferred box check) against the oracle's all-boxes loop."""
    m = soup(po, seed)
    s = HostSimScene(hostsim, rtp, m, with_boxes=True)
    views, lights, dirl = SOUP_VIEWS, SOUP_LIGHTS, SOUP_DIRL
    for view in views:
        a = s.render(40, 36, view, 3, lights, dirl)
        b = po.render(m, 40, 36, view=view, max_depth=3, point_lights=lights, dir_lights=dirl)
        assert np.array_equal(a["hit"], b["hit"]), f"{(a['hit'] != b['hit']).sum()} primary ids differ"
        assert bits_equal(a["t"], b["t"]).all()
        nan = np.isnan(b["rgb"])
        assert np.array_equal(np.isnan(a["rgb"]), nan)
        assert bits_equal(np.where(nan, 0, a["rgb"]), np.where(nan, 0, b["rgb"])).all()
        assert a["counts"].tolist() == [b["counts"][k] for k in ("primary", "shadow", "reflection", "hits")]
    s.close()
