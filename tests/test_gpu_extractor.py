"""GPU parity: CUDA ORBextractor (through the C ABI) vs the CPU oracle -- bit-exact keypoints + descriptors."""
import numpy as np
import pytest

from gf_orb_slam_b200 import synth

pytestmark = pytest.mark.gpu


def _cmp_frame(orc, ex_cpu, img, b, kps, desc, counts):
    k, d = ex_cpu(img)
    n = counts[b]
    assert n == len(k), (n, len(k))
    gk, gd = kps[b, :n], desc[b, :n]
    for f in ("octave", "x", "y", "size", "response", "class_id"):
        assert np.array_equal(gk[f], k[f]), f
    # orientation: north_star asks 1e-5 relative; the path is un-fused IEEE so it is in fact bit-exact
    assert np.array_equal(gk["angle"], k["angle"])
    assert np.array_equal(gd, d)


def _same_outputs(a, b):
    """(keypoints, descriptors, counts) of two runs agree byte for byte on the rows that were written (rows beyond counts[b]
    are never written by gforb_extract and hold whatever the allocation held)"""
    assert np.array_equal(a[2], b[2])
    for f in range(len(a[2])):
        n = int(a[2][f])
        assert a[0][f, :n].tobytes() == b[0][f, :n].tobytes(), f
        assert np.array_equal(a[1][f, :n], b[1][f, :n]), f


@pytest.mark.parametrize("w,h,nf", [(640, 480, 1000), (752, 480, 1000), (1241, 376, 2000), (200, 152, 300)])
def test_extract_matches_oracle(orc, w, h, nf):
    from gf_orb_slam_b200 import ORBextractor
    B = 3
    imgs = np.stack([synth.textured_frame(w, h, seed=10 + i, band=(i != 1)) for i in range(B)])
    ex = ORBextractor(nf, 1.2, 8, max_width=w, max_height=h, max_batch=B)
    kps, desc, counts = ex.extract_batch(imgs)
    ex_cpu = orc.ORBextractor(nf, 1.2, 8)
    # pyramid taps of frame 0 (ComputePyramid output, before any blur)
    lv = ex_cpu.pyramid(imgs[0])
    for l in range(8):
        assert np.array_equal(ex.debug_level(0, l, 0), lv[l]), f"pyramid level {l}"
        assert np.array_equal(ex.debug_level(0, l, 1)[16:-16, 16:-16], orc.blur(lv[l][16:-16, 16:-16], 0)), f"blur level {l}"
    for b in range(B):
        _cmp_frame(orc, ex_cpu, imgs[b], b, kps, desc, counts)
    assert ex.last_launches() > 0


@pytest.mark.parametrize("w,h,nf,variant", [(640, 480, 1000, 0), (752, 480, 500, 1), (200, 152, 300, 0)])
def test_harris_score_matches_oracle(orc, w, h, nf, variant):
    """scoreType = HARRIS_SCORE (src/ORBextractor.cc:86-127, :667-671): float responses drive both retainBest passes"""
    from gf_orb_slam_b200 import HARRIS_SCORE, ORBextractor
    B = 3
    imgs = np.stack([synth.textured_frame(w, h, seed=40 + i, band=(i != 1)) for i in range(B)])
    ex = ORBextractor(nf, 1.2, 8, HARRIS_SCORE, max_width=w, max_height=h, max_batch=B, retain_variant=variant)
    kps, desc, counts = ex.extract_batch(imgs)
    ex_cpu = orc.ORBextractor(nf, 1.2, 8, scoreType=0, retainVariant=variant)
    for b in range(B):
        _cmp_frame(orc, ex_cpu, imgs[b], b, kps, desc, counts)
    k, _ = ex_cpu(imgs[0])
    assert len(np.unique(k["response"])) > 200 and (k["response"] != np.round(k["response"])).any()   # really Harris floats
    # and the FAST_SCORE keypoint set differs (the mode is not a no-op)
    kf, _ = orc.ORBextractor(nf, 1.2, 8, scoreType=1, retainVariant=variant)(imgs[0])
    assert len(kf) != len(k) or not (np.array_equal(kf["x"], k["x"]) and np.array_equal(kf["y"], k["y"]))


def test_sequence_frames_and_single_call(orc):
    from gf_orb_slam_b200 import ORBextractor
    seq = synth.Sequence()
    imgs = seq.frames(0, 4)
    ex = ORBextractor(1000, 1.2, 8, max_batch=4)
    kps, desc, counts = ex.extract_batch(imgs)
    ex_cpu = orc.ORBextractor(1000, 1.2, 8)
    for b in range(4):
        _cmp_frame(orc, ex_cpu, imgs[b], b, kps, desc, counts)
    k1, d1 = ex(imgs[2])                      # operator() call shape, batch of one
    assert np.array_equal(k1, kps[2, :counts[2]]) and np.array_equal(d1, desc[2, :counts[2]])
    k0, d0 = ex(np.zeros((0, 0), np.uint8))   # empty image -> silent empty return (src/ORBextractor.cc:772)
    assert len(k0) == 0 and d0.shape == (0, 32)


def test_edge_images(orc):
    from gf_orb_slam_b200 import ORBextractor
    w, h = 640, 480
    flat = np.full((h, w), 77, np.uint8)                       # no corners anywhere
    weak = synth.textured_frame(w, h, seed=5)
    weak = (weak.astype(np.int32) // 8 + 100).astype(np.uint8)  # low contrast: mostly the minThFAST fallback
    rng = np.random.default_rng(0)
    noise = rng.integers(0, 256, (h, w), dtype=np.uint8)       # maximum corner density / ties
    imgs = np.stack([flat, weak, noise])
    ex = ORBextractor(1000, 1.2, 8, max_width=w, max_height=h, max_batch=3)
    kps, desc, counts = ex.extract_batch(imgs)
    ex_cpu = orc.ORBextractor(1000, 1.2, 8)
    assert counts[0] == 0
    for b in range(3):
        _cmp_frame(orc, ex_cpu, imgs[b], b, kps, desc, counts)


@pytest.mark.parametrize("blur_variant,retain_variant", [(1, 0), (0, 1)])
def test_variants(orc, blur_variant, retain_variant):
    from gf_orb_slam_b200 import ORBextractor
    img = synth.textured_frame(640, 480, seed=21)
    ex = ORBextractor(1000, 1.2, 8, blur_variant=blur_variant, retain_variant=retain_variant, max_width=640, max_height=480)
    kps, desc, counts = ex.extract_batch(img[None])
    ex_cpu = orc.ORBextractor(1000, 1.2, 8, blurVariant=blur_variant, retainVariant=retain_variant)
    _cmp_frame(orc, ex_cpu, img, 0, kps, desc, counts)


@pytest.mark.parametrize("size,nf", [((1241, 376), 2000), ((752, 480), 1000), ((200, 152), 500)])
def test_resize_variants_agree(monkeypatch, size, nf):
    """the pyramid resize with source rows staged by cp.async.bulk (GFORB_RESIZE=tma, the A/B partner recorded in
    profiles/r2_ncu_summary.txt) and with aligned word loads + dp2a (GFORB_RESIZE=words / words8: k_resize2 with 4- / 8-row bands) write the same pyramid
    bytes -- ring and pitch padding columns included -- and the same keypoints / descriptors as the byte-gather kernel"""
    from gf_orb_slam_b200 import ORBextractor
    w, h = size
    imgs = np.stack([synth.textured_frame(w, h, seed=40 + i) for i in range(2)])
    res = []
    for v in ("direct", "tma", "words", "words8"):
        monkeypatch.setenv("GFORB_RESIZE", v)
        ex = ORBextractor(nf, 1.2, 8, max_width=w, max_height=h, max_batch=2)
        kps, desc, counts = ex.extract_batch(imgs)
        res.append((kps.copy(), desc.copy(), counts.copy(), [ex.debug_level(1, l, 0).copy() for l in range(8)]))
    a = res[0]
    for b in res[1:]:
        _same_outputs(a, b)
        for l in range(8):
            assert np.array_equal(a[3][l], b[3][l]), l


@pytest.mark.parametrize("size,nf", [((752, 480), 1000), ((1241, 376), 2000), ((200, 152), 1500)])
@pytest.mark.parametrize("knob", ["GFORB_KP=v1", "GFORB_DESCRIBE=smem", "GFORB_DESCRIBE=magic", "GFORB_BLUR=v2", "GFORB_BLUR_QD=1", "GFORB_BLUR_QD=3", "GFORB_FN_TALL=0", "GFORB_FN_TALL=1"])
def test_kernel_generations_agree(monkeypatch, knob, size, nf):
    """A/B partners of the extractor kernels against the default build, byte for byte (keypoints incl. angles, descriptors,
    counts, blurred planes): GFORB_KP=v1 (first-generation k_orient / slot-mapped k_describe instead of k_orient2 and the
    row-mapped k_describe), GFORB_DESCRIBE=smem (k_describe2: tap window staged in shared memory, cvRound by the 1.5 * 2^23
    add; the A/B loser), GFORB_BLUR=v2 (vertical blur pass on unpacked lane-ops instead of k_blur3's IDP.2A row pairs),
    GFORB_FN_TALL=0/1 (strips of <= 64 / <= 128 rows in k_fast_nms and the blur; the default picks by batch size).
    The small frame puts most keypoints within 18 px of the ROI edge, so k_describe2's blurred / unblurred-ring merge runs."""
    from gf_orb_slam_b200 import ORBextractor
    w, h = size
    imgs = np.stack([synth.textured_frame(w, h, seed=70 + i) for i in range(3)])
    res = []
    for v in (None, knob):
        for k in ("GFORB_KP", "GFORB_DESCRIBE", "GFORB_BLUR", "GFORB_BLUR_QD", "GFORB_FN_TALL"):
            monkeypatch.delenv(k, raising=False)
        if v:
            monkeypatch.setenv(*v.split("="))
        ex = ORBextractor(nf, 1.2, 8, max_width=w, max_height=h, max_batch=3)
        kps, desc, counts = ex.extract_batch(imgs)
        res.append((kps.copy(), desc.copy(), counts.copy(), [ex.debug_level(2, l, 1)[16:-16, 16:-16].copy() for l in range(8)]))
    a, b = res
    _same_outputs(a, b)
    for l in range(8):
        assert np.array_equal(a[3][l], b[3][l]), l
    assert a[2].min() > 50


@pytest.mark.parametrize("size,nf,variant", [((752, 480), 1000, 0), ((1241, 376), 2000, 1), ((640, 480), 300, 0)])
def test_retain_variants_agree(monkeypatch, size, nf, variant):
    """retainBest's nth_element replayed by the 32 lanes of a warp (default) gives the permutation of the sequential replay
    (GFORB_RETAIN=seq, the libstdc++ introselect of csrc/introselect.h): same keypoints in the same order, same descriptors.
    Small nfeatures make every cell and level overflow its quota, so the selection runs everywhere."""
    from gf_orb_slam_b200 import ORBextractor
    w, h = size
    imgs = np.stack([synth.textured_frame(w, h, seed=60 + i) for i in range(3)])
    res = []
    for v in ("seq", "warp"):
        monkeypatch.setenv("GFORB_RETAIN", v)
        ex = ORBextractor(nf, 1.2, 8, max_width=w, max_height=h, max_batch=3, retain_variant=variant)
        kps, desc, counts = ex.extract_batch(imgs)
        res.append((kps.copy(), desc.copy(), counts.copy()))
    a, b = res
    _same_outputs(a, b)
    assert a[2].min() > nf // 2


@pytest.mark.parametrize("w,h,nf", [(752, 480, 1000), (640, 480, 1000), (1241, 376, 2000), (320, 240, 500)])
def test_quadtree_distribution_matches_oracle(orc, w, h, nf):
    """the non-default ORB-SLAM2 quadtree distribution (gforb_extractor_set_distribution(h, 1)) against the restatement in
    oracle/orb_quadtree.cpp: keypoints (order included), angles and descriptors bit-exact; the mode differs from the default;
    switching back gives the default mode's result again"""
    from gf_orb_slam_b200 import ORBextractor
    B = 3
    imgs = np.stack([synth.textured_frame(w, h, seed=80 + i, band=(i != 1)) for i in range(B)])
    imgs[2, :, : w // 2] = 90                                             # a frame with a flat half: few keys, early termination
    ex = ORBextractor(nf, 1.2, 8, max_width=w, max_height=h, max_batch=B)
    k0, d0, c0 = [a.copy() for a in ex.extract_batch(imgs)]
    ex.set_distribution(1)
    kps, desc, counts = [a.copy() for a in ex.extract_batch(imgs)]
    ex_cpu = orc.ORBextractor(nf, 1.2, 8)
    ex_cpu.set_distribution(1)
    for b in range(B):
        _cmp_frame(orc, ex_cpu, imgs[b], b, kps, desc, counts)
    assert counts[0] > nf // 2 and counts.max() <= nf
    assert not (np.array_equal(counts, c0) and kps[0, :c0[0]].tobytes() == k0[0, :c0[0]].tobytes())
    ex.set_distribution(0)
    k1, d1, c1 = ex.extract_batch(imgs)
    assert np.array_equal(c1, c0)
    for b in range(B):                                                    # rows past counts[b] are not written
        assert k1[b, :c0[b]].tobytes() == k0[b, :c0[b]].tobytes() and np.array_equal(d1[b, :c0[b]], d0[b, :c0[b]])


def test_quadtree_distribution_rejects_degenerate_grids():
    """a pyramid whose top level is lower than one 30-px cell: the published code divides by zero there (the oracle returns an
    error), the product answers GFORB_ERR_GEOMETRY and stays usable in the default mode"""
    from gf_orb_slam_b200 import ORBextractor
    from gf_orb_slam_b200._lib import GforbError
    ex = ORBextractor(300, 1.2, 8, max_width=200, max_height=152, max_batch=1)
    with pytest.raises(GforbError):
        ex.set_distribution(1)
    ex.set_distribution(0)
    img = synth.textured_frame(200, 152, seed=3)
    kps, desc, counts = ex.extract_batch(img[None])
    assert counts[0] > 100
