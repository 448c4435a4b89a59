"""CPU-side checks of the drop-in boundary: libgforb.so loads, exports every symbol include/gforb.h declares, the
header and the C++ shim compile as plain C / C++17, and argument validation works without a device."""
import ctypes
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, "include", "gforb.h")
LIB = os.path.join(ROOT, "gf_orb_slam_b200", "libgforb.so")


def declared_symbols():
    src = open(HDR).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gforb_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_path():
    syms = declared_symbols()
    for need in ("gforb_extract", "gforb_descriptor_distance", "gforb_search_by_projection", "gforb_search_for_initialization",
                 "gforb_obs_build_info_frame", "gforb_obs_select", "gforb_allpairs_top2", "gforb_top2_merge"):
        assert need in syms


def test_library_exports_every_declared_symbol():
    assert os.path.exists(LIB), "libgforb.so missing: run python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(LIB)
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing
    lib.gforb_version.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.gforb_version()


def test_python_binding_table_covers_header():
    from gf_orb_slam_b200 import _lib
    lib = _lib.lib()
    for s in declared_symbols():
        assert getattr(lib, s) is not None


def test_header_is_plain_c_and_shim_is_cxx17(tmp_path):
    c = tmp_path / "t.c"
    c.write_text('#include "gforb.h"\nint main(void){ GforbKeyPoint k; (void)k; return sizeof(GforbKeyPoint) == 28 ? 0 : 1; }\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(c), "-o", str(tmp_path / "t")])
    subprocess.check_call([str(tmp_path / "t")])
    cc = tmp_path / "s.cpp"
    cc.write_text('#include "gforb/orb_slam.hpp"\nint main(){ return sizeof(gforb_cv::KeyPoint) == 28 ? 0 : 1; }\n')
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(cc), "-o", str(tmp_path / "s.o")])


def test_shim_opencv_branch_compiles_against_the_opencv_interface(tmp_path):
    """include/gforb/orb_slam.hpp with -DGFORB_WITH_OPENCV (the branch a maintainer of the reference builds): every class
    instantiated and every member the reference's call sites use named with cv::Mat / cv::KeyPoint / cv::Point2f arguments,
    compiled against a stand-in for the OpenCV headers (tests/cpp/opencv_stub; OpenCV's C++ headers are not in this image) and
    linked against libgforb.so, so that every C-ABI symbol the branch references resolves."""
    src = os.path.join(ROOT, "tests", "cpp", "shim_opencv_branch.cpp")
    out = tmp_path / "shim_cv"
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           "-I", os.path.join(ROOT, "tests", "cpp", "opencv_stub"), src, "-o", str(out),
                           "-L", os.path.join(ROOT, "gf_orb_slam_b200"), "-lgforb", "-L/usr/local/cuda/lib64",
                           "-Wl,-rpath," + os.path.join(ROOT, "gf_orb_slam_b200"), "-Wl,-rpath,/usr/local/cuda/lib64"])
    subprocess.check_call([str(out)])          # main() touches nothing that needs a device


def test_no_device_means_error_not_fallback():
    """Without a CUDA device every create() must fail loudly (GFORB_ERR_CUDA); bad arguments are rejected first."""
    from gf_orb_slam_b200 import _lib
    lib = _lib.lib()
    h = ctypes.c_void_p()
    assert lib.gforb_extractor_create(None, ctypes.byref(h)) == -2
    cfg = _lib.ExtractorConfig(nfeatures=1000, scaleFactor=1.2, nlevels=8, scoreType=1, iniThFAST=20, minThFAST=7,
                               blur_variant=0, retain_variant=0, max_width=640, max_height=480, max_batch=1, device=0)
    rc = lib.gforb_extractor_create(ctypes.byref(cfg), ctypes.byref(h))
    if lib.gforb_device_count() == 0:
        assert rc == -1 and not h.value
    else:
        assert rc == 0
        lib.gforb_extractor_destroy(h)
    cfg.scoreType = 2          # only HARRIS_SCORE (0) and FAST_SCORE (1) exist: argument error, never a silent substitute
    assert lib.gforb_extractor_create(ctypes.byref(cfg), ctypes.byref(h)) == -2
