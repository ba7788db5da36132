"""rtp_trace_rays_device_noalloc (include/rtp.h): the ray-batch launches with their scratch memory (ray cursor,
parked-ray counters and slots) in a static __device__ array of the library instead of the scene's cudaMalloc'ed
blocks.  No counterpart in the reference.  Checked through the C ABI against rtp_trace_rays_device, bit for bit
(face ids, t, occlusion bytes - integer / index work and correctly rounded float chains: no tolerance), on two calls
in a row per scene: the array keeps the previous call's contents, the second call must not see them.
Collected last on purpose."""
import numpy as np
import pytest

from conftest import load_golden_mesh


def _rays(rng, n, span):
    o = rng.uniform(-span, span, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    return np.concatenate([o, d], 1)


def _compare(rtp, torch, s, rays_np):
    dev = torch.device("cuda:0")
    n = len(rays_np)
    rays = torch.from_numpy(rays_np).to(dev)
    face = [torch.full((n,), -7, dtype=torch.int32, device=dev) for _ in range(2)]
    t = [torch.full((n,), -7.0, dtype=torch.float32, device=dev) for _ in range(2)]
    occ = [torch.full((n,), 9, dtype=torch.uint8, device=dev) for _ in range(2)]
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    s.trace_rays_device(rays.data_ptr(), n, rtp.RAYS_CLOSEST, d_face=face[0].data_ptr(), d_t=t[0].data_ptr(), stream=stream)
    s.trace_rays_device(rays.data_ptr(), n, rtp.RAYS_ANYHIT, d_occ=occ[0].data_ptr(), stream=stream)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    s.trace_rays_device_noalloc(rays.data_ptr(), n, rtp.RAYS_CLOSEST, d_face=face[1].data_ptr(), d_t=t[1].data_ptr(), stream=stream)
    s.trace_rays_device_noalloc(rays.data_ptr(), n, rtp.RAYS_ANYHIT, d_occ=occ[1].data_ptr(), stream=stream)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before, "the no-alloc entry point changed the device's free memory"
    assert torch.equal(face[0], face[1]), "closest-hit face ids differ"
    assert torch.equal(t[0].view(torch.int32), t[1].view(torch.int32)), "closest-hit t differs"
    assert torch.equal(occ[0], occ[1]), "any-hit bytes differ"
    hits = int((face[1] >= 0).sum())
    assert 0 < hits < n and int(occ[1].max()) == 1 and int(occ[1].min()) == 0  # the batch exercised hits and misses
    return hits


@pytest.mark.gpu
def test_noalloc_ray_batches_equal_the_allocating_entry_point(rtp):
    import torch
    from importlib import import_module
    synth = import_module("ray-tracing-project_b200.synth")
    rng = np.random.default_rng(2024)
    # (a) a scene larger than the L2 (1 M triangles): closest-hit batches park their stragglers, so the static
    #     parked-ray slots and k_coop's cursor are exercised; the 200 k-ray batch overflows the 4 096 slots
    #     (a ray that finds the list full carries on), the 3 k-ray one does not
    m = synth.terrain(708)
    s = rtp.Scene(m["verts"], m["normals"], m["faces"], m["face_mat"], m["materials"], m["shape"], device=0)
    s.set_lights(None, None)
    _compare(rtp, torch, s, _rays(rng, 200_000, 0.6))
    _compare(rtp, torch, s, _rays(rng, 3_001, 0.6))  # second call in a row: shorter batch, stale scratch behind it
    s.close()
    # (b) an L2-resident scene (no parking), a different scene handle on the same device: the array is per device
    g = load_golden_mesh("dodgeColorTest")
    s = rtp.Scene(g.verts, g.normals, g.faces, g.face_mat, g.materials, g.shape, device=0)
    s.set_lights(None, None)
    _compare(rtp, torch, s, _rays(rng, 50_000, 1.2))
    _compare(rtp, torch, s, _rays(rng, 50_000, 1.2))
    s.close()


@pytest.mark.gpu
def test_noalloc_limits_are_enforced(rtp):
    import ctypes as C
    import torch
    g = load_golden_mesh("dodgeColorTest")
    s = rtp.Scene(g.verts, g.normals, g.faces, g.face_mat, g.materials, g.shape, device=0)
    L = rtp.lib()
    buf = torch.zeros(64, dtype=torch.float32, device="cuda:0")
    too_many = (1 << 22) + 1
    rc = L.rtp_trace_rays_device_noalloc(s._h, buf.data_ptr(), too_many, rtp.RAYS_ANYHIT, None, None, buf.data_ptr(), None)
    assert rc == -1 and b"exceed the limit" in L.rtp_last_error()
    rc = L.rtp_trace_rays_device_noalloc(s._h, buf.data_ptr(), 4, rtp.RAYS_SHADE, None, None, buf.data_ptr(), None)
    assert rc == -1 and b"mode" in L.rtp_last_error()
    rc = L.rtp_trace_rays_device_noalloc(s._h, buf.data_ptr(), 4, rtp.RAYS_CLOSEST, None, None, None, None)
    assert rc == -1 and b"face output" in L.rtp_last_error()
    assert L.rtp_trace_rays_device_noalloc(s._h, buf.data_ptr(), 0, rtp.RAYS_ANYHIT, None, None, buf.data_ptr(), None) == 0
    s.close()
    assert C.sizeof(C.c_void_p) == 8
