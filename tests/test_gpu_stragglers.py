"""The straggler path of the trace kernels (k_trace parks long rays, k_coop finishes each one with a
whole warp - csrc/rtp_kernels.cu) must give the same hits as the serial traversal.  The policy is
read from the environment when the library first launches a kernel, so the parity suite is re-run
in a child process with parking forced on for EVERY ray after its first step."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("env", [
    dict(RTP_BUDGET="1", RTP_PARK_LANES="32", RTP_PARK_ANY="1", RTP_PARK_SMALL="1"),                  # everything parks
    dict(RTP_BUDGET="2", RTP_PARK_LANES="8", RTP_PARK_ANY="1", RTP_PARK_SMALL="1", RTP_REFILL="24"),  # mixed, lane refill
])
def test_parity_suite_with_forced_parking(env):
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_parity.py"),
                        os.path.join(ROOT, "tests", "test_gpu_workloads.py"), "-m", "gpu", "-x", "-q"],
                       env=e, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
