"""Multi-device scenes driven from ONE process (rtp_scene_create_multi; Flyscene::DEVICES; `raytrace --gpus N`):
the frame is cut into interleaved screen tiles, every rank renders its share on its own stream (fed by its own host
thread) and stores the finished pixels into the first device's image; fork and join are CUDA events.
Replaces the std::thread fan-out + join of the reference's raytraceScene (src/flyscene.cpp:266-289).

On a one-GPU box the ranks share the GPU (devices = [0, 0, ...]) - same code path, the peer stores just stay on the
device; with more GPUs visible the same tests spread the ranks over them (NVLink peer stores)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, bits_equal, load_case, load_golden_mesh

pytestmark = pytest.mark.gpu
CLI = os.path.join(ROOT, "ray-tracing-project_b200", "raytrace")
REF_MODELS = os.path.join(ROOT, "oracle", "_ref", "resources", "models")


def scene_on(rtp, m, devices):
    return rtp.Scene(m.verts, m.normals, m.faces, m.face_mat, m.materials, m.shape, boxes=m.boxes, box_off=m.box_off,
                     box_faces=m.box_faces, devices=devices)


def rank_devices(rtp, n):
    ndev = rtp.device_count()
    return [r % ndev for r in range(n)]


@pytest.mark.parametrize("n", [2, 3, 8])
@pytest.mark.parametrize("case,depth", [("dodge_c2", 4), ("cube_dirlight", 2)])
def test_multi_device_frame_is_bitwise_the_single_device_frame(rtp, n, case, depth):
    c = load_case(case)
    m = load_golden_mesh(str(c["model"]))
    w, h = 333, 190  # ragged right / bottom tiles
    dirl = c["dir_lights"] if "dir_lights" in c and len(c["dir_lights"]) else None
    frames = []
    for devices in ([0], rank_devices(rtp, n)):
        s = scene_on(rtp, m, devices)
        assert s.device_count() == len(devices)
        s.set_lights(c["point_lights"], dirl)
        s.set_camera(w, h, c["view"])
        frames.append(s.render(w, h, max_depth=depth, want=("rgb", "rgb8", "hit")))
        if len(devices) > 1:  # a second frame through the same handle (buffers, workers and events are reused)
            again = s.render(w, h, max_depth=depth, want=("rgb8", "hit"))
            assert np.array_equal(again["rgb8"], frames[-1]["rgb8"]) and np.array_equal(again["hit"], frames[-1]["hit"])
        s.close()
    a, b = frames
    assert np.array_equal(a["hit"], b["hit"])
    assert bits_equal(a["rgb"], b["rgb"]).all()
    assert np.array_equal(a["rgb8"], b["rgb8"])
    for k in ("rays_primary", "rays_shadow", "rays_reflection", "hits"):
        assert a["stats"][k] == b["stats"][k], k  # the ranks' counters add up to the one-device frame's
    assert (a["hit"] >= 0).any()


def test_multi_device_scene_rejects_external_sharding(rtp):
    c = load_case("cube_rot")
    m = load_golden_mesh(str(c["model"]))
    s = scene_on(rtp, m, rank_devices(rtp, 2))
    s.set_lights(c["point_lights"], None)
    s.set_camera(64, 64, c["view"])
    with pytest.raises(rtp.RtpError):
        s.render(64, 64, rank=0, world=2)
    s.close()


def test_cli_gpus_flag_writes_the_same_ppm(tmp_path):
    """`raytrace --gpus N` (Flyscene::DEVICES) writes the result.ppm `--gpus 1` writes, byte for byte."""
    if not os.path.exists(CLI):
        pytest.skip("raytrace CLI not built")
    model = os.path.join(REF_MODELS, "dodgeColorTest.obj")
    if not os.path.exists(model):
        pytest.skip("shipped models not present")
    c = load_case("dodge_c2")
    outs = []
    for n in (1, 8):
        out = str(tmp_path / f"r{n}.ppm")
        cmd = [CLI, "--model", model, "--size", "256", "144", "--depth", "3", "--out", out, "--gpus", str(n), "--nolights",
               "--view"] + [repr(float(x)) for x in c["view"]]
        for l in c["point_lights"]:
            cmd += ["--light"] + [repr(float(x)) for x in l]
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        outs.append(open(out, "rb").read())
    assert outs[0] == outs[1] and len(outs[0]) > 256 * 144 * 6
