"""Parity tests proper: the CUDA kernels, called through the C ABI (include/rtp.h), against the
reference goldens and the oracle on the same inputs.

Bar (BASELINE.json north_star): primary-hit face ids identical except documented edge cases;
per-pixel RGB within 1/255 on >= 99.9 % of pixels.  What is actually asserted is stricter:
face ids and primary t bit-identical on every pixel, RGB within 1/255 on every non-NaN pixel
(the only non-bit-exact step is CUDA powf vs glibc powf in the specular term)."""
import numpy as np
import pytest

from conftest import bits_equal, golden_cases, load_case, load_golden_mesh, soup, SOUP_VIEWS, SOUP_LIGHTS, SOUP_DIRL

pytestmark = pytest.mark.gpu
TOL = 1.0 / 255.0


def make_scene(rtp, m, with_boxes=False, build=0):
    kw = dict(boxes=m.boxes, box_off=m.box_off, box_faces=m.box_faces) if with_boxes else {}
    return rtp.Scene(m.verts, m.normals, m.faces, m.face_mat, m.materials, m.shape, build=build, **kw)


def check_image(got_rgb, got_hit, ref_rgb, ref_hit, exact_fraction=0.5):
    assert np.array_equal(got_hit, ref_hit), f"{(got_hit != ref_hit).sum()} primary-hit ids differ"
    nan_ref, nan_got = np.isnan(ref_rgb), np.isnan(got_rgb)
    assert np.array_equal(nan_ref, nan_got), "NaN pixels (SURVEY Q8) differ"
    diff = np.abs(np.where(nan_ref, 0, got_rgb) - np.where(nan_ref, 0, ref_rgb))
    assert diff.max() <= TOL, f"max RGB error {diff.max()} > 1/255"
    assert bits_equal(got_rgb, ref_rgb).all(-1).mean() >= exact_fraction


@pytest.mark.parametrize("name", golden_cases())
def test_render_matches_reference_golden(rtp, name):
    c = load_case(name)
    m = load_golden_mesh(c["model"])
    w, h = int(c["w"]), int(c["h"])
    s = make_scene(rtp, m)
    s.set_lights(c["point_lights"], c["dir_lights"])
    s.set_camera(w, h, c["view"])
    out = s.render(w, h, max_depth=int(c["max_depth"]), want=("rgb", "rgb8", "hit"))
    s.close()
    check_image(out["rgb"].reshape(-1, 3), out["hit"].ravel(), c["rgb"], c["hit"])
    st = out["stats"]
    assert [st["rays_primary"], st["rays_shadow"], st["rays_reflection"], st["hits"]] == c["counts"].tolist()
    # quantised image = what writePPMImage would print (ppmIO.hpp:150) wherever RGB is bit-equal
    import pyoracle
    q = pyoracle.quantise(c["rgb"])
    same = bits_equal(out["rgb"].reshape(-1, 3), c["rgb"]).all(-1)
    assert np.array_equal(out["rgb8"].reshape(-1, 3)[same], q[same])
    assert (np.abs(out["rgb8"].reshape(-1, 3).astype(int) - q.astype(int)) <= 1).all()


def test_full_hd_frame_against_oracle_subsample(rtp, po):
    """BASELINE config 2 shape (1920x1080, 4 lights, max_depth 4) on the shipped default scene;
    the oracle checks a pixel subsample, size-independent properties cover the rest."""
    c = load_case("dodge_c2")
    m = load_golden_mesh("dodgeColorTest")
    w, h = 1920, 1080
    s = make_scene(rtp, m)
    s.set_lights(c["point_lights"], None)
    s.set_camera(w, h, c["view"])
    out = s.render(w, h, max_depth=4, want=("rgb", "hit"))
    px = po.all_pixels(w, h, stride=(13, 11), offset=(3, 5))
    o = po.render(m, w, h, view=c["view"], max_depth=4, point_lights=c["point_lights"], pixels=px, threads=8)
    got_rgb, got_hit = out["rgb"][px[:, 1], px[:, 0]], out["hit"][px[:, 1], px[:, 0]]
    check_image(got_rgb, got_hit, o["rgb"], o["hit"])
    # properties at full size: background exactly where nothing was hit, colours in [0, 1]
    miss = out["hit"] < 0
    assert (out["rgb"][miss] == np.float32(0.9)).all()
    assert (out["rgb"][~miss] >= 0).all() and (out["rgb"][~miss] <= 1).all()
    assert out["stats"]["rays_primary"] == w * h and out["stats"]["hits"] >= (~miss).sum()
    # idempotence: a second render of the same frame is bit-identical
    out2 = s.render(w, h, max_depth=4, want=("rgb", "hit"))
    s.close()
    assert np.array_equal(out["hit"], out2["hit"]) and bits_equal(out["rgb"], out2["rgb"]).all()


def test_ragged_image_sizes(rtp, po):
    """Sizes that are not multiples of the 32x32 tile / 8x4 warp footprint, and W != H (Q1)."""
    m = load_golden_mesh("testding")
    s = make_scene(rtp, m)
    s.set_lights(None, None)
    for (w, h) in [(1, 1), (7, 3), (33, 31), (100, 37), (37, 100)]:
        s.set_camera(w, h, None)
        out = s.render(w, h, max_depth=2)
        o = po.render(m, w, h)
        check_image(out["rgb"].reshape(-1, 3), out["hit"].ravel(), o["rgb"], o["hit"], exact_fraction=0.3)
    s.close()


def test_tile_sharding_is_bitwise_equal_to_one_gpu(rtp):
    """Every rank's interleaved tiles, rendered separately, assemble into the 1-GPU image."""
    c = load_case("dodge_c2")
    m = load_golden_mesh("dodgeColorTest")
    w, h = 200, 150
    s = make_scene(rtp, m)
    s.set_lights(c["point_lights"], None)
    s.set_camera(w, h, c["view"])
    full = s.render(w, h, max_depth=3)
    for world in (2, 3, 8):
        acc = dict(rgb=np.full((h, w, 3), -7, np.float32), hit=np.full((h, w), -7, np.int32))
        rays = 0
        for r in range(world):
            part = s.render(w, h, max_depth=3, rank=r, world=world, out=acc)
            rays += part["stats"]["rays_primary"]
        assert rays == w * h
        assert np.array_equal(acc["hit"], full["hit"]) and bits_equal(acc["rgb"], full["rgb"]).all()
    s.close()


def test_fused_resolve_writes_the_same_image_as_gather_and_untile(rtp):
    """rtp_render_image_device: every (virtual) rank resolves its tiles straight into one shared image -
    the multi-GPU drivers point it at GPU 0's memory - and the result must equal the un-tile path bit
    for bit: quantised rgb8 (ppmIO.hpp:150) and primary face ids, including ragged edge tiles."""
    import torch
    c = load_case("dodge_c2")
    m = load_golden_mesh("dodgeColorTest")
    w, h = 203, 149  # not multiples of the 32-pixel tile
    s = make_scene(rtp, m)
    s.set_lights(c["point_lights"], None)
    s.set_camera(w, h, c["view"])
    full = s.render(w, h, max_depth=3, want=("rgb", "rgb8", "hit"))           # un-tile path (float plane requested)
    fused1 = s.render(w, h, max_depth=3, want=("rgb8", "hit"))                # fused path inside rtp_render
    assert np.array_equal(fused1["rgb8"], full["rgb8"]) and np.array_equal(fused1["hit"], full["hit"])
    st = torch.cuda.current_stream().cuda_stream
    for world in (1, 2, 5, 8):
        img = torch.full((h, w, 3), 77, dtype=torch.uint8, device="cuda")
        hit = torch.full((h, w), -9, dtype=torch.int32, device="cuda")
        for r in range(world):
            s.render_image_device(w, h, d_rgb8=img.data_ptr(), d_hit=hit.data_ptr(), max_depth=3, rank=r, world=world, stream=st)
        torch.cuda.synchronize()
        assert np.array_equal(img.cpu().numpy(), full["rgb8"]), f"rgb8 differs at world {world}"
        assert np.array_equal(hit.cpu().numpy(), full["hit"]), f"face ids differ at world {world}"
    s.close()


@pytest.mark.parametrize("name", ["dodge_c2", "dodge_c1", "testding_2l", "glas_nan", "quirks"])
def test_gpu_built_bvh_gives_the_same_image(rtp, name):
    """RTP_BUILD_GPU_LBVH (csrc/bvh_build_gpu.cu: Morton sort, radix tree, bottom-up fit on the device) only
    changes how candidates are culled: hits, t order and colours must equal the host-SAH scene bit for bit."""
    c = load_case(name)
    m = load_golden_mesh(c["model"])
    w, h = int(c["w"]), int(c["h"])
    out = []
    for build in (rtp.BUILD_HOST_SAH, rtp.BUILD_GPU_LBVH):
        s = make_scene(rtp, m, build=build)
        s.set_lights(c["point_lights"], c["dir_lights"])
        s.set_camera(w, h, c["view"])
        out.append(s.render(w, h, max_depth=int(c["max_depth"])))
        info = s.info()
        assert info["n_triangles"] == len(m.faces) and info["n_bvh_nodes"] > 0
        s.close()
    assert np.array_equal(out[0]["hit"], out[1]["hit"])
    assert bits_equal(out[0]["rgb"], out[1]["rgb"]).all()
    check_image(out[1]["rgb"].reshape(-1, 3), out[1]["hit"].ravel(), c["rgb"], c["hit"])


def test_reference_box_filter_is_reproduced(rtp, po):
    """Faces whose BoundingBox fails the reference's own slab test are invisible (Q11/Q13): with
    artificially shrunken boxes this hits the deferred visibility check and the exact re-run."""
    from test_hostsim import shrunken_boxes_mesh
    m2 = shrunken_boxes_mesh(po)
    c = load_case("dodge_c2")
    w, h = 320, 180
    s = make_scene(rtp, m2, with_boxes=True)
    s.set_lights(c["point_lights"], None)
    s.set_camera(w, h, c["view"])
    out = s.render(w, h, max_depth=3)
    s.close()
    o = po.render(m2, w, h, view=c["view"], max_depth=3, point_lights=c["point_lights"], threads=8)
    check_image(out["rgb"].reshape(-1, 3), out["hit"].ravel(), o["rgb"], o["hit"])


def test_ray_batch_api_matches_oracle(rtp, po):
    m = load_golden_mesh("dodgeColorTest")
    rng = np.random.default_rng(11)
    n = 20000
    o = rng.uniform(-1.2, 1.2, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    d[:50, 0] = 0.0
    d[50:100, 1] = 0.0
    d[100:120, :2] = 0.0
    rays = np.concatenate([o, d], 1)
    s = make_scene(rtp, m)
    s.set_lights(None, None)
    g = s.trace_rays(rays, rtp.RAYS_CLOSEST)
    f2, t2 = po.closest_rays(m, rays, threads=8)
    assert np.array_equal(g["face"], f2) and bits_equal(g["t"], t2).all()
    g = s.trace_rays(rays, rtp.RAYS_ANYHIT)
    assert np.array_equal(g["occluded"].astype(np.int32), po.shadow_rays(m, rays, threads=8))
    sub = rays[:2000]
    g = s.trace_rays(sub, rtp.RAYS_SHADE, max_depth=3)
    ref = po.trace_rays(m, sub, max_depth=3)
    s.close()
    nan = np.isnan(ref)
    assert np.array_equal(nan, np.isnan(g["rgb"]))
    assert np.abs(np.where(nan, 0, g["rgb"]) - np.where(nan, 0, ref)).max() <= TOL


def test_error_paths(rtp):
    m = load_golden_mesh("plane")
    s = make_scene(rtp, m)
    with pytest.raises(rtp.RtpError):  # camera / lights not set yet
        s.render(8, 8)
    s.set_lights(None, None)
    s.set_camera(8, 8)
    with pytest.raises(rtp.RtpError):
        s.render(8, 8, max_depth=9)
    with pytest.raises(rtp.RtpError):
        s.render(0, 8)
    s.close()
    bad = m.faces.copy()
    bad[0, 0] = 99
    with pytest.raises(rtp.RtpError):
        rtp.Scene(m.verts, m.normals, bad, m.face_mat, m.materials, m.shape)


@pytest.mark.parametrize("seed", range(8))
def test_tie_heavy_random_scenes_match_the_oracle(rtp, po, seed):
    """The fuzz of tests/test_hostsim.py on the device: triangle soups with coplanar, duplicated and degenerate
    faces seen along grid-aligned rays - primary ids identical, RGB within 1/255 (bit-equal except powf)."""
    m = soup(po, seed, n_tris=400)
    s = make_scene(rtp, m, with_boxes=True)
    s.set_lights(SOUP_LIGHTS, SOUP_DIRL)
    for view in SOUP_VIEWS:
        s.set_camera(72, 56, view)
        out = s.render(72, 56, max_depth=3)
        o = po.render(m, 72, 56, view=view, max_depth=3, point_lights=SOUP_LIGHTS, dir_lights=SOUP_DIRL)
        check_image(out["rgb"].reshape(-1, 3), out["hit"].ravel(), o["rgb"], o["hit"], exact_fraction=0.5)
        st = out["stats"]
        assert [st["rays_primary"], st["rays_shadow"], st["rays_reflection"], st["hits"]] == \
               [o["counts"][k] for k in ("primary", "shadow", "reflection", "hits")]
    s.close()


def test_shadow_ray_count_beyond_32_bits_is_refused(rtp):
    """shadow rays of a level are indexed (pixel, light) in 32 bits.  A frame whose pixels x lights would
    overflow must be refused with an error, not traced with a wrapped count."""
    m = load_golden_mesh("plane")
    s = make_scene(rtp, m)
    lights = np.tile(np.float32([[0.1, 2.0, 3.0, 0.001, 0.001, 0.001]]), (1100, 1))
    lights[:, 0] += np.arange(1100, dtype=np.float32) * 1e-3
    s.set_lights(lights, None)
    s.set_camera(1920, 1080)
    with pytest.raises(rtp.RtpError, match="shadow rays"):      # 2 088 960 tile-padded pixels x 1100 lights > 2^31
        s.render(1920, 1080, max_depth=1, want=("rgb8",))
    s.set_camera(64, 64)
    out = s.render(64, 64, max_depth=1, want=("hit",))            # the same 1100 lights on a small frame are fine
    assert out["stats"]["rays_shadow"] == out["stats"]["hits"] * 1100 and out["stats"]["hits"] > 0
    s.close()


@pytest.mark.parametrize("seed,dist", [(0, 300.0), (1, 3000.0), (2, 40000.0)])
def test_far_camera_matches_the_oracle(rtp, po, seed, dist):
    """P = o + t*d carries ulp(|o|) of rounding, which grows with the camera distance while the build-time
    padding of the hierarchy does not.  rtp_set_camera widens the boxes for far cameras (ensure_pad); primary ids and
    colours must still be the oracle's, which tests every face."""
    m = soup(po, 100 + seed, n_tris=300)
    s = make_scene(rtp, m, with_boxes=True)
    s.set_lights(SOUP_LIGHTS, None)
    w, h = 64, 48
    fovy = float(np.degrees(2 * np.arctan(1.6 / dist)))          # narrow lens: the scene still fills the frame
    for axis in range(3):
        view = np.zeros((3, 4), np.float32)
        view[:, :3] = np.roll(np.eye(3, dtype=np.float32), axis, axis=1)
        view[2, 3] = -dist
        s.set_camera(w, h, view.ravel(), fovy=fovy)
        out = s.render(w, h, max_depth=2)
        o = po.render(m, w, h, view=view.ravel(), fovy=fovy, max_depth=2, point_lights=SOUP_LIGHTS)
        assert (o["hit"] >= 0).mean() > 0.05
        check_image(out["rgb"].reshape(-1, 3), out["hit"].ravel(), o["rgb"], o["hit"], exact_fraction=0.5)
    s.close()


@pytest.mark.parametrize("name", ["dodge_c2", "cube_rot", "plane_diag", "quirks", "soft17_dodge"])
def test_path_kernel_equals_the_per_level_kernels(rtp, name, monkeypatch):
    """Bounce levels >= 1 run either path by path in one persistent kernel (k_paths: RTP_DEEP=paths, the default for
    small frames) or as per-level closest / shadow / shade launches (RTP_DEEP=wavefront, the default for big frames);
    the variable is read when the scene is created.  Same functions, same arithmetic:
    the two frames must agree bit for bit, ray counters included."""
    c = load_case(name)
    m = load_golden_mesh(c["model"])
    w, h = 3 * int(c["w"]), 3 * int(c["h"])
    frames = []
    for mode in ("paths", "wavefront"):
        monkeypatch.setenv("RTP_DEEP", mode)
        s = make_scene(rtp, m)
        s.set_lights(c["point_lights"], c["dir_lights"])
        s.set_camera(w, h, c["view"])
        frames.append(s.render(w, h, max_depth=6, want=("rgb", "hit")))
        s.close()
    a, b = frames
    assert np.array_equal(a["hit"], b["hit"]) and bits_equal(a["rgb"], b["rgb"]).all()
    for k in ("rays_primary", "rays_shadow", "rays_reflection", "hits"):
        assert a["stats"][k] == b["stats"][k], k
    assert a["stats"]["kernel_launches"] < b["stats"]["kernel_launches"]


@pytest.mark.parametrize("name", ["dodge_c2", "cube_rot", "plane_diag", "quirks", "soft17_dodge", "glas_nan"])
@pytest.mark.parametrize("depth", [1, 4])
def test_camera_kernel_equals_the_separate_launches(rtp, name, depth, monkeypatch):
    """The camera level runs in one kernel (k_camera: primary rays, their shadow rays, shading and - max_depth 1 - the
    image; RTP_FUSED=1, the default) or as separate closest / shadow / shade launches (RTP_FUSED=0); read when the scene
    is created.  Same functions, same arithmetic: every output must agree bit for bit, ray counters included - also
    with every ray forced through the parked-pixel kernel (k_coop_camera)."""
    c = load_case(name)
    m = load_golden_mesh(c["model"])
    w, h = 3 * int(c["w"]) + 5, 3 * int(c["h"]) + 3  # ragged: tile padding pixels
    frames = []
    for fused in ("0", "1"):
        monkeypatch.setenv("RTP_FUSED", fused)
        s = make_scene(rtp, m)
        s.set_lights(c["point_lights"], c["dir_lights"])
        s.set_camera(w, h, c["view"])
        frames.append(s.render(w, h, max_depth=depth, want=("rgb", "rgb8", "hit")))
        s.close()
    a, b = frames
    assert np.array_equal(a["hit"], b["hit"]) and bits_equal(a["rgb"], b["rgb"]).all() and np.array_equal(a["rgb8"], b["rgb8"])
    for k in ("rays_primary", "rays_shadow", "rays_reflection", "hits"):
        assert a["stats"][k] == b["stats"][k], k
    assert b["stats"]["kernel_launches"] < a["stats"]["kernel_launches"]


@pytest.mark.parametrize("fused", ["0", "1"])
@pytest.mark.parametrize("depth", [1, 3])
def test_frame_without_lights_and_with_directional_lights_only(rtp, po, fused, depth, monkeypatch):
    """Edge cases of the light loop on both camera-level paths: no light at all (every hit shades to black, reflections
    still bounce: flyscene.cpp:603-614 sums nothing), and directional lights only (flyscene.cpp:611)."""
    monkeypatch.setenv("RTP_FUSED", fused)
    c = load_case("cube_dirlight")
    m = load_golden_mesh(c["model"])
    w, h = 2 * int(c["w"]) + 7, 2 * int(c["h"]) + 1
    none = np.zeros((0, 6), np.float32)  # (None would mean the reference's default light)
    for point, dirl in ((none, none), (none, c["dir_lights"])):
        s = make_scene(rtp, m)
        s.set_lights(point, dirl)
        s.set_camera(w, h, c["view"])
        out = s.render(w, h, max_depth=depth)
        s.close()
        o = po.render(m, w, h, view=c["view"], max_depth=depth, point_lights=point, dir_lights=dirl, threads=4)
        check_image(out["rgb"].reshape(-1, 3), out["hit"].ravel(), o["rgb"], o["hit"])
        assert out["stats"]["rays_shadow"] == o["counts"]["shadow"] and out["stats"]["rays_reflection"] == o["counts"]["reflection"]
        if len(dirl) == 0:
            assert out["stats"]["rays_shadow"] == 0 and out["stats"]["hits"] > 0
