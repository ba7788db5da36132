"""rtp_render_async / rtp_render_wait (asynchronous, double-buffered device-to-host output for
multi-frame renders): every frame of a camera path must land in its host buffers exactly as the synchronous
rtp_render delivers it, with the same ray counters, while two frames are in flight."""
import numpy as np
import pytest

from conftest import load_case, load_golden_mesh

pytestmark = pytest.mark.gpu


def orbit(synth, k, n):
    a = 2 * np.pi * k / n
    return synth.look_at((2.0 * np.sin(a), 0.4 + 0.1 * k, 2.0 * np.cos(a)), (0.0, 0.0, 0.0))


def test_camera_path_async_equals_sync(rtp):
    import importlib
    import torch
    synth = importlib.import_module("ray-tracing-project_b200.synth")
    c = load_case("dodge_c2")
    m = load_golden_mesh("dodgeColorTest")
    w, h, depth, n = 400, 232, 3, 7
    s = rtp.Scene(m.verts, m.normals, m.faces, m.face_mat, m.materials, m.shape)
    s.set_lights(c["point_lights"], None)
    s.set_tile_groups(3)
    want = []
    for k in range(n):
        s.set_camera(w, h, orbit(synth, k, n))
        out = s.render(w, h, max_depth=depth, want=("rgb8", "hit"))
        want.append((out["rgb8"].copy(), out["hit"].copy(), out["stats"]))
    assert not np.array_equal(want[0][0], want[1][0])
    rgb8 = [torch.zeros(h, w, 3, dtype=torch.uint8).pin_memory() for _ in range(2)]
    hit = [torch.zeros(h, w, dtype=torch.int32).pin_memory() for _ in range(2)]
    got = []

    def collect(k):
        st = s.render_wait()
        got.append((rgb8[k & 1].numpy().copy(), hit[k & 1].numpy().copy(), st))

    for k in range(n):
        if k >= 2:
            collect(k - 2)  # frees host slot k & 1
        s.set_camera(w, h, orbit(synth, k, n))
        s.render_async(w, h, rgb8[k & 1].data_ptr(), hit[k & 1].data_ptr(), max_depth=depth)
    collect(n - 2)
    collect(n - 1)
    for k in range(n):
        assert np.array_equal(got[k][0], want[k][0]), f"frame {k}: image differs"
        assert np.array_equal(got[k][1], want[k][1]), f"frame {k}: primary ids differ"
        for key in ("rays_primary", "rays_shadow", "rays_reflection", "hits", "kernel_launches"):
            assert got[k][2][key] == want[k][2][key], (k, key)
        assert got[k][2]["ms_trace"] > 0
    s.close()


def test_async_call_order_is_checked(rtp):
    import torch
    c = load_case("plane_c1")
    m = load_golden_mesh("plane")
    s = rtp.Scene(m.verts, m.normals, m.faces, m.face_mat, m.materials, m.shape)
    s.set_lights(c["point_lights"], None)
    s.set_camera(64, 64, c["view"])
    buf = [torch.zeros(64, 64, 3, dtype=torch.uint8).pin_memory() for _ in range(3)]
    with pytest.raises(rtp.RtpError, match="no frame outstanding"):
        s.render_wait()
    with pytest.raises(rtp.RtpError, match="no output plane"):
        s.render_async(64, 64)
    s.render_async(64, 64, buf[0].data_ptr())
    s.render_async(64, 64, buf[1].data_ptr())
    with pytest.raises(rtp.RtpError, match="already outstanding"):
        s.render_async(64, 64, buf[2].data_ptr())
    with pytest.raises(rtp.RtpError, match="outstanding"):
        s.render(64, 64, want=("rgb8",))
    s.render_wait()
    s.render_wait()
    assert np.array_equal(buf[0].numpy(), buf[1].numpy())
    out = s.render(64, 64, want=("rgb8",))
    assert np.array_equal(out["rgb8"], buf[0].numpy())
    # pageable host memory works too (the copies then block the caller instead of overlapping)
    plain, plain_hit = np.zeros((64, 64, 3), np.uint8), np.zeros((64, 64), np.int32)
    s.render_async(64, 64, plain.ctypes.data, plain_hit.ctypes.data)
    st = s.render_wait()
    assert np.array_equal(plain, out["rgb8"]) and st["rays_primary"] == 64 * 64
    assert np.array_equal(plain_hit, s.render(64, 64, want=("hit",))["hit"])
    s.close()


def test_peer_barrier_kernel_arrives_waits_and_times_out(rtp):
    """rtp_peer_barrier on one GPU: a satisfied target passes at once, the counter only grows, and a rank that never
    arrives costs the others a bounded wait (counted), not a hang."""
    import time
    import torch
    c = load_case("plane_c1")
    m = load_golden_mesh("plane")
    s = rtp.Scene(m.verts, m.normals, m.faces, m.face_mat, m.materials, m.shape)
    counter = torch.zeros(64, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    for frame in (1, 2, 3):                       # world = 1: the rank's own arrival satisfies the target
        s.peer_barrier(counter.data_ptr(), frame, 500, stream)
    torch.cuda.synchronize()
    assert int(counter[0].item()) == 3 and s.peer_barrier_timeouts() == 0
    t0 = time.perf_counter()
    s.peer_barrier(counter.data_ptr(), 2 * 4, 50, stream)   # "world = 2", frame 4: the other rank never comes
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    assert 0.04 < dt < 1.0, dt
    assert int(counter[0].item()) == 4 and s.peer_barrier_timeouts() == 1
    with pytest.raises(rtp.RtpError):
        s.peer_barrier(counter.data_ptr() + 2, 1, 50, stream)  # misaligned counter
    s.close()
