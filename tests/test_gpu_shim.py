"""The C++ drop-in shim (include/gforb/orb_slam.hpp: ORB_SLAM::ORBextractor / ORBmatcher / Observability over the C ABI)
against the CPU oracle, as one native test binary (tests/cpp/test_shim.cpp, built by __graft_entry__.build())."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "test_shim.bin")


@pytest.mark.gpu
def test_cpp_shim_matches_oracle():
    if not os.path.exists(BIN):
        import __graft_entry__ as g
        g.build_shim_test()
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    print(r.stdout[-2000:], r.stderr[-2000:])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "OK" in r.stdout
