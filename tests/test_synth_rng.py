"""The counter-based generator behind BASELINE config 5 (ray-tracing-project_b200/synth.py): Philox4x32-10 against
the published known-answer vectors (Random123 kat_vectors), and the properties C5 needs - rays keyed by pixel index,
identical from numpy (tests) and torch (bench.py), uniform on the hemisphere.  CPU only."""
import importlib

import numpy as np

synth = importlib.import_module("ray-tracing-project_b200.synth")


def kat(counter, key):
    r = synth.philox4x32(*[np.array([x], np.int64) for x in counter], key[0], key[1])
    return [int(v[0]) for v in r]


def test_philox4x32_known_answers():
    assert kat((0, 0, 0, 0), (0, 0)) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert kat((0xffffffff,) * 4, (0xffffffff,) * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert kat((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0)) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_ao_directions_are_keyed_by_pixel_and_library_independent():
    import torch
    rng = np.random.default_rng(0)
    N = rng.normal(size=(300, 3)).astype(np.float32)
    N /= np.linalg.norm(N, axis=1, keepdims=True)
    pix = rng.choice(1920 * 1080, 300, replace=False).astype(np.int64)
    P = rng.normal(size=(300, 3)).astype(np.float32)
    a = synth.ao_rays_pixels(P, N, pix, 64)
    b = synth.ao_rays_pixels(torch.from_numpy(P), torch.from_numpy(N), torch.from_numpy(pix), 64, xp=torch).numpy()
    assert a.dtype == np.float32 and a.shape == (300 * 64, 6)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    perm = rng.permutation(300)
    c = synth.ao_rays_pixels(P[perm], N[perm], pix[perm], 64)
    assert np.array_equal(c.reshape(300, 64, 6), a.reshape(300, 64, 6)[perm])
    assert not np.array_equal(synth.ao_rays_pixels(P, N, pix, 64, seed=1235), a)
    d = a.reshape(300, 64, 6)[:, :, 3:]
    assert np.abs(np.linalg.norm(d, axis=2) - 1).max() < 1e-6
    cos = (d * N[:, None, :]).sum(2)
    assert (cos >= 0).all() and abs(cos.mean() - 0.5) < 0.01          # uniform on the hemisphere: E[cos] = 1/2
    assert np.array_equal(a.reshape(300, 64, 6)[:, 0, :3], (P + np.float32(0.003) * N).astype(np.float32))
