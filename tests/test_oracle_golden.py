"""The oracle (oracle/oracle.c) against the known answers generated from the reference itself
(tests/golden/*.npz, made by tests/golden/make_golden.py from oracle/_ref/ref_harness)."""
import numpy as np
import pytest

from conftest import bits_equal, golden_cases, load_case, load_golden_mesh


@pytest.mark.parametrize("name", golden_cases())
def test_oracle_matches_reference_golden(po, name):
    c = load_case(name)
    m = load_golden_mesh(c["model"])
    o = po.render(m, int(c["w"]), int(c["h"]), view=c["view"], max_depth=int(c["max_depth"]),
                  point_lights=c["point_lights"], dir_lights=c["dir_lights"])
    assert bits_equal(o["dir"], c["dir"]).all(), "primary directions differ"
    assert np.array_equal(o["hit"], c["hit"]), "primary-hit face ids differ"
    assert bits_equal(o["t"], c["t"]).all(), "primary t differs"
    assert bits_equal(o["rgb"], c["rgb"]).all(), "RGB differs (bit-exact expected: same libm)"
    assert [o["counts"][k] for k in ("primary", "shadow", "reflection", "hits")] == c["counts"].tolist()


@pytest.mark.parametrize("model", ["plane", "cube", "testding", "glas", "dodgeColorTest"])
def test_oracle_mesh_prep_matches_reference(po, model):
    """createFaces normals, computeNormals, normalizeModelMatrix and the BoundingBox splitter."""
    m = load_golden_mesh(model)
    m2 = po.Mesh(m.verts, None if model in ("cube", "testding", "glas") else m.normals, m.faces, m.face_mat,
                 m.materials).build_boxes()
    assert int(np.bitwise_xor.reduce(m2.face_normal.view(np.uint32).ravel())) == m.face_normal_crc
    assert np.array_equal(m2.shape.view(np.int32), m.shape.view(np.int32))
    assert m2.boxes.shape == m.boxes.shape and np.array_equal(m2.boxes.view(np.int32), m.boxes.view(np.int32))
    assert np.array_equal(m2.box_off, m.box_off) and np.array_equal(m2.box_faces, m.box_faces)
    if model in ("cube", "testding", "glas"):  # #vn != #v in the OBJ: smooth normals recomputed
        assert np.array_equal(m2.normals.view(np.int32), m.normals.view(np.int32))


def test_oracle_camera_matches_reference(po):
    for name in golden_cases():
        c = load_case(name)
        eye, inv = po.camera_derive(int(c["w"]), int(c["h"]), c["view"])
        cam = c["camera"]
        assert np.array_equal(eye.view(np.int32), cam[:3].view(np.int32))
        assert np.array_equal(inv.view(np.int32), cam[15:27].view(np.int32))


def test_oracle_threads_agree(po):
    c = load_case("dodge_c2")
    m = load_golden_mesh(c["model"])
    a = po.render(m, 48, 27, view=c["view"], max_depth=3, point_lights=c["point_lights"], threads=1)
    b = po.render(m, 48, 27, view=c["view"], max_depth=3, point_lights=c["point_lights"], threads=4)
    assert bits_equal(a["rgb"], b["rgb"]).all() and np.array_equal(a["hit"], b["hit"])
    assert a["counts"] == b["counts"]


def test_quantiser(po):
    x = np.array([0.0, 0.5, 1.0, 0.999, 0.00392, 0.00393, np.nan], np.float32)
    q = po.quantise(x)
    assert q.tolist() == [0, 127, 255, 254, 0, 1, 0]
