"""The drop-in boundary: libflyscene_b200.so loads and exports every symbol include/rtp.h
declares; without a CUDA device it fails loudly instead of falling back to the CPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, has_cuda, load_golden_mesh


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "rtp.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rtp_[a-z_0-9]+)\s*\(", src)))


def test_header_symbols_are_exported(rtp):
    L = rtp.lib()
    names = declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/rtp.h but not exported"
    assert sorted(rtp.EXPORTS) == names
    assert L.rtp_abi_version() == 7


def test_struct_sizes_match_header(rtp):
    # plain C structs: sizes follow from the header's field lists (LP64)
    assert C.sizeof(rtp.RtpMesh) == 16 + 6 * 8 + 48 + 3 * 8 + 16
    assert C.sizeof(rtp.RtpCamera) == 48 + 16 + 8
    assert C.sizeof(rtp.RtpRenderOpts) == 12 + 12 + 12
    assert C.sizeof(rtp.RtpStats) == 32 + 8 + 12 + 12 + 8


def test_invalid_arguments_are_rejected(rtp):
    L = rtp.lib()
    assert L.rtp_scene_create(None, 0, None) == -1
    assert b"null" in L.rtp_last_error()
    assert L.rtp_tile_buffer_bytes(0, 10, 0, 1) == 0
    assert L.rtp_tile_buffer_bytes(64, 64, 0, 1) == 4 * 1024 * 16
    assert L.rtp_tile_buffer_bytes(1920, 1080, 0, 8) == 256 * 1024 * 16  # 60 x 34 = 2040 tiles over 8 ranks, padded to the largest share


@pytest.mark.skipif(has_cuda(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback_without_a_gpu(rtp):
    m = load_golden_mesh("plane")
    with pytest.raises(rtp.RtpError) as e:
        rtp.Scene(m.verts, m.normals, m.faces, m.face_mat, m.materials, m.shape)
    assert "no CUDA device" in str(e.value) or "rtp error -2" in str(e.value)


def test_product_does_not_reference_the_oracle():
    """oracle/ is test infrastructure: nothing under the package may import, include or link it."""
    pkg = os.path.join(ROOT, "ray-tracing-project_b200")
    for d, _, files in os.walk(pkg):
        if os.sep + "build" in d:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".hpp", "Makefile")):
                txt = open(os.path.join(d, f), errors="replace").read()
                assert "pyoracle" not in txt and "liboracle" not in txt and "oracle/" not in txt, os.path.join(d, f)
    out = os.popen(f"ldd {os.path.join(pkg, 'libflyscene_b200.so')}").read()
    assert "oracle" not in out and "hostsim" not in out
