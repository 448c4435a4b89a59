"""Known-answer tests of the oracle's per-keypoint arithmetic (SURVEY 8(c) "new KATs"): the reference's tests hold no
vectors for extraction, so these are derived by hand / by an independent numpy restatement.
  * steered BRIEF at the four cardinal angles: the rotated tap of pattern point (x, y) is exactly (x, y), (-y, x),
    (-x, -y), (y, -x), so the descriptor is a pure table lookup on the patch (src/ORBextractor.cc:161-201)
  * IC_Angle on patches whose intensity centroid is known in closed form (src/ORBextractor.cc:131-158)
  * hand-checkable 64x64 frames: a single bright pixel is the only FAST-9 corner; a bright square's corner pixels are
    corners before NMS and all lose the strict-> NMS against their equal-score neighbours"""
import os
import re

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(os.path.dirname(HERE), "gf_orb_slam_b200", "csrc")


@pytest.fixture(scope="module")
def orc():
    import oracle
    oracle.build()
    return oracle


def _pattern():
    txt = open(os.path.join(CSRC, "brief_pattern.inc")).read()
    v = [tuple(int(t) for t in m) for m in re.findall(r"\{(-?\d+),(-?\d+),(-?\d+),(-?\d+)\}", txt)]
    assert len(v) == 256
    return np.array(v, np.int32)


def _descriptor_lookup(img, cx, cy, rot):
    """bit k of byte i = I(tap0) < I(tap1) for pattern pair 8 i + k, taps rotated by rot * 90 degrees"""
    P = _pattern()
    out = np.zeros(32, np.uint8)

    def tap(x, y):
        for _ in range(rot):                      # (x, y) -> (x cos - y sin, x sin + y cos) at 90 degrees = (-y, x)
            x, y = -y, x
        return int(img[cy + y, cx + x])
    for i in range(32):
        v = 0
        for k in range(8):
            x0, y0, x1, y1 = P[i * 8 + k]
            v |= int(tap(x0, y0) < tap(x1, y1)) << k
        out[i] = v
    return out


@pytest.mark.parametrize("rot", [0, 1, 2, 3])
def test_descriptor_at_cardinal_angles(orc, rot):
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (64, 64), dtype=np.uint8)
    got = orc.describe_at(img, 32.0, 31.0, 90.0 * rot)
    exp = _descriptor_lookup(img, 32, 31, rot)
    assert np.array_equal(got, exp)
    # and the four descriptors differ from each other (the steering is not a no-op)
    if rot:
        assert not np.array_equal(got, orc.describe_at(img, 32.0, 31.0, 0.0))


def test_descriptor_of_flat_and_ramp_patch(orc):
    flat = np.full((64, 64), 90, np.uint8)
    assert not orc.describe_at(flat, 32.0, 32.0, 37.0).any()                  # I(t0) < I(t1) never holds
    ramp = np.tile(np.arange(64, dtype=np.uint8) * 3, (64, 1))                # intensity grows with x
    P = _pattern()
    exp = np.zeros(32, np.uint8)
    for i in range(32):
        for k in range(8):
            exp[i] |= int(P[i * 8 + k][0] < P[i * 8 + k][2]) << k             # x0 < x1
    assert np.array_equal(orc.describe_at(ramp, 32.0, 32.0, 0.0), exp)


def test_ic_angle_closed_form(orc):
    yy, xx = np.mgrid[0:64, 0:64]
    right = np.where(xx > 32, 200, 10).astype(np.uint8)                       # mass to the right: m01 = 0, m10 > 0
    assert orc.ic_angle_at(right, 32.0, 32.0) == 0.0
    down = np.where(yy > 32, 200, 10).astype(np.uint8)                        # mass below: m10 = 0, m01 > 0 -> 90 degrees
    assert abs(orc.ic_angle_at(down, 32.0, 32.0) - 90.0) < 1e-3
    left = np.where(xx < 32, 200, 10).astype(np.uint8)
    assert abs(orc.ic_angle_at(left, 32.0, 32.0) - 180.0) < 1e-3
    # a single bright pixel at offset (u, v) = (5, -12): angle = fastAtan2(-12, 5), fastAtan2 is accurate to 0.3 degrees
    one = np.zeros((64, 64), np.uint8); one[32 - 12, 32 + 5] = 255
    exp = np.degrees(np.arctan2(-12.0, 5.0)) % 360.0
    assert abs(orc.ic_angle_at(one, 32.0, 32.0) - exp) < 0.3
    # exact integer moments against a direct sum over the circular patch (749 pixels)
    rng = np.random.default_rng(9)
    img = rng.integers(0, 256, (64, 64), dtype=np.uint8)
    umax = [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]         # src/ORBextractor.cc:497-522 for HALF_PATCH_SIZE 15
    m10 = m01 = n = 0
    for v in range(-15, 16):
        for u in range(-umax[abs(v)], umax[abs(v)] + 1):
            m10 += u * int(img[32 + v, 32 + u]); m01 += v * int(img[32 + v, 32 + u]); n += 1
    assert n == 749
    assert orc.ic_angle_at(img, 32.0, 32.0) == pytest.approx(float(orc.fast_atan2(np.float32([m01]), np.float32([m10]))[0]), abs=0)


def test_hand_checkable_frame(orc):
    """64x64 frames whose FAST-9 answer follows by hand: a single bright pixel is the only corner (its 16 ring pixels are all
    darker, every other pixel sees at most one brighter ring pixel), its score is the contrast minus one; the pixels of a bright
    square that are corners without NMS are those within the corner regions, and NMS removes them all (equal scores, strict >)"""
    img = np.zeros((64, 64), np.uint8)
    img[30, 25] = 180
    k = orc.fast(img, 20, True)
    assert [(int(p["x"]), int(p["y"]), int(p["response"])) for p in k] == [(25, 30, 179)]
    assert len(orc.fast(img, 180, True)) == 0                       # threshold is strict: |diff| must exceed it
    sq = np.zeros((64, 64), np.uint8)
    sq[22:42, 22:42] = 200
    raw = orc.fast(sq, 20, False)
    pts = {(int(p["x"]), int(p["y"])) for p in raw}
    assert {(22, 22), (22, 41), (41, 22), (41, 41)} <= pts and (32, 32) not in pts and (22, 32) not in pts
    assert len(orc.fast(sq, 20, True)) == 0


def test_quadtree_oracle_properties(orc):
    """oracle/orb_quadtree.cpp (ORB-SLAM2's distribution, the non-default mode; the reference holds no such code and no vectors):
    every level returns at most mnFeaturesPerLevel + 3 keypoints, the frame at most nfeatures, no pixel twice, and on a
    frame whose right half is flat every keypoint comes from the textured half -- while a frame with a single corner returns it"""
    from gf_orb_slam_b200 import synth
    img = synth.textured_frame(752, 480, seed=11)
    ex = orc.ORBextractor(1000, 1.2, 8)
    ex.set_distribution(1)
    k, d = ex(img)
    per = np.bincount(k["octave"], minlength=8)
    fpl = np.array(ex.features_per_level())
    assert len(k) <= 1000 and (per <= fpl + 3).all() and per[0] >= fpl[0]
    assert len({(float(a), float(b), int(o)) for a, b, o in zip(k["x"], k["y"], k["octave"])}) == len(k)
    half = img.copy(); half[:, 376:] = 77
    kh, _ = ex(half)
    assert len(kh) > 200 and (kh["x"] < 376 + 4 * 1.2 ** kh["octave"]).all()
    one = np.full((480, 752), 30, np.uint8); one[200, 300] = 220
    k1, _ = ex(one)
    assert len(k1) >= 1 and (k1["x"][0], k1["y"][0], k1["octave"][0]) == (300.0, 200.0, 0)
