"""The parity suite and the launch-sequence workload on the bounds-checked build (libflyscene_b200_checked.so, -DRTP_CHECKED:
csrc/rtp_device.h) - the stand-in for compute-sanitizer memcheck, which is closed on this pool.  pytest runs in THIS process,
so the library instance whose check block is read at the end is the one the tests used.
usage: RTP_LIB=ray-tracing-project_b200/libflyscene_b200_checked.so python profiles/checked_run.py [pytest args]"""
import ctypes as C
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("RTP_LIB", os.path.join(ROOT, "ray-tracing-project_b200", "libflyscene_b200_checked.so"))
import pytest  # noqa: E402

rtp = importlib.import_module("ray-tracing-project_b200")
L = rtp.lib()
assert hasattr(L, "rtp_debug_checks"), "not the checked build"
args = sys.argv[1:] or [os.path.join(ROOT, "tests", f) for f in ("test_gpu_parity.py", "test_gpu_workloads.py", "test_gpu_groups.py",
                                                                  "test_gpu_async.py", "test_gpu_multi_device.py", "test_zz_no_alloc.py")]
rc = pytest.main(args + ["-m", "gpu", "-x", "-q", "-p", "no:cacheprovider"])
out = (C.c_uint * 4)()
assert L.rtp_debug_checks(out) == 0
env = {k: v for k, v in os.environ.items() if k.startswith("RTP_") and k != "RTP_LIB"}
print(f"checked build, env {env}: pytest rc {int(rc)}, failed bounds checks {out[0]} (first: code {out[1]}, value {out[2]}, limit {out[3]})")
sys.exit(int(rc) or (1 if out[0] else 0))
