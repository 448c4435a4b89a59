"""Known-answer tests of the constant tables the product ships (they are #included by the oracle too, so a wrong table
would pass every GPU-vs-oracle parity test): the rBRIEF pattern against SURVEY.md App. A.9 and -- when the reference
tree is present (this container; never on the GPU box) -- against src/ORBextractor.cc:204-462 itself; glibc's log()
table against the running libm; the device's restated log()/sinf()/cosf() host builds against libm (tests/cpp)."""
import hashlib
import os
import re
import struct
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "gf_orb_slam_b200", "csrc")
BRIEF_SHA256 = "7e645581387b82784797e8adddb9b6f0c12611859fda09ca8a9bec96d767a05f"     # begins 7e645581387b8278


def _brief_ints():
    txt = open(os.path.join(CSRC, "brief_pattern.inc")).read()
    txt = re.sub(r"//.*", "", txt)
    return [int(x) for x in re.findall(r"-?\d+", txt)]


def test_brief_pattern_sha256_and_kats():
    nums = _brief_ints()
    assert len(nums) == 1024
    assert hashlib.sha256(struct.pack("<1024i", *nums)).hexdigest() == BRIEF_SHA256
    pts = [(nums[2 * i], nums[2 * i + 1]) for i in range(512)]
    assert pts[:4] == [(8, -3), (9, 5), (4, 2), (7, -12)] and pts[-2:] == [(-1, -6), (0, -11)]
    assert max(abs(x) for x, _ in pts) == 13 and max(abs(y) for _, y in pts) == 13
    assert max(x * x + y * y for x, y in pts) == 338                                                # (-13,-13): radius 18.385


def test_brief_pattern_equals_reference_source():
    ref = "/root/reference/src/ORBextractor.cc"
    if not os.path.exists(ref):
        pytest.skip("reference tree not present (GPU box)")
    src = open(ref).read().split("\n")
    txt = "\n".join(src[203:462])
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    txt = re.sub(r"//.*", "", txt)
    body = txt[txt.index("{") + 1: txt.rindex("}")]
    assert [int(x) for x in re.findall(r"-?\d+", body)] == _brief_ints()


def test_glibc_log_table_matches_running_libm():
    """csrc/glibc_log_data.inc was read out of libm.so.6 (tools/gen_glibc_log_table.py); a different libm on this host
    would make the oracle's log() and the device's restatement disagree -- fail here, loudly, instead."""
    libm = "/lib/x86_64-linux-gnu/libm.so.6"
    if not os.path.exists(libm):
        pytest.skip("no x86-64 glibc libm at the usual path")
    words = [int(x, 16) for x in re.findall(r"0x([0-9a-f]{16})ull", open(os.path.join(CSRC, "glibc_log_data.inc")).read())]
    assert len(words) == 18 + 256
    assert struct.pack("<%dQ" % len(words), *words) in open(libm, "rb").read()


@pytest.mark.parametrize("src,args", [("test_glibc_log.cpp", ["12"]), ("test_sincosf.cpp", ["97"]), ("test_introselect.cpp", [])])
def test_restated_libm_and_libstdcxx_routines(tmp_path, src, args):
    exe = str(tmp_path / "t.bin")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-pthread", os.path.join(ROOT, "tests", "cpp", src), "-o", exe])
    out = subprocess.run([exe] + args, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
