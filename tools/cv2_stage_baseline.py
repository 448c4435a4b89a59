"""the OpenCV calls the reference's extractor makes (copyMakeBorder, the resize chain, FAST per cell with
the threshold fallback, GaussianBlur 7x7) issued through cv2 (SIMD / IPP builds), one frame per worker PROCESS (the
python binding builds KeyPoint lists under the GIL, so threads would serialise).  Orientation, descriptors, matching and
selection are the reference's own scalar C++ and are NOT in this figure: it bounds how much the scalar oracle port
understates the OpenCV-backed stages.  Stand-alone on purpose (no CUDA, no torch): bench.py runs it as a subprocess.
usage: python tools/cv2_stage_baseline.py <frames> <workers>   -> one JSON line"""
import importlib.util
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W_, H_ = 752, 480


def _load(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "gf_orb_slam_b200", name + ".py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def one(img):
    import cv2
    cv2.setNumThreads(1)
    sizes = one.sizes
    nfeat = [217, 181, 151, 126, 105, 87, 73, 60]                     # features per level for 1000
    det20 = cv2.FastFeatureDetector_create(20, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    det7 = cv2.FastFeatureDetector_create(7, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    lv = cv2.copyMakeBorder(img, 16, 16, 16, 16, cv2.BORDER_REFLECT_101)
    n = 0
    for l, (w, h) in enumerate(sizes):
        if l:
            roi = cv2.resize(lv[16:-16, 16:-16], (w, h), interpolation=cv2.INTER_LINEAR)
            lv = cv2.copyMakeBorder(roi, 16, 16, 16, 16, cv2.BORDER_REFLECT_101)
        Wb, Hb = w - 32, h - 32                                       # maxBorder - minBorder (src/ORBextractor.cc:544-560)
        cols = max(int(np.sqrt(np.float32(nfeat[l]) / np.float32(5.0 * Wb / Hb))), 1)
        rows = max(int(np.float32(Wb / Hb) * cols), 1)
        cw, ch = -(-Wb // cols), -(-Hb // rows)
        for i in range(rows):
            for j in range(cols):
                y0, x0 = 32 + i * ch - 3, 32 + j * cw - 3             # plane coordinates: ring 16 + border 16
                y1, x1 = min(y0 + ch + 6, 32 + Hb + 3), min(x0 + cw + 6, 32 + Wb + 3)
                if y1 <= y0 or x1 <= x0:
                    continue
                cell = lv[y0:y1, x0:x1]
                k = det20.detect(cell)
                if len(k) <= 3:
                    k = det7.detect(cell)
                n += len(k)
        cv2.GaussianBlur(lv[16:-16, 16:-16], (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
    return n


def main():
    nfr = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    workers = int(sys.argv[2]) if len(sys.argv) > 2 else (os.cpu_count() or 1)
    try:
        import cv2
    except Exception as e:                                             # no OpenCV in the image: say so, measure nothing
        print(json.dumps({"unavailable": "cv2 not importable: %s" % e}))
        return
    synth = _load("synth")
    consts = _load("pipeline_consts")
    one.sizes = consts.level_sizes(W_, H_)
    frames = list(synth.Sequence(W_, H_).frames(0, nfr))
    with mp.get_context("fork").Pool(workers) as pool:
        pool.map(one, frames[:workers])                                # warm-up: imports, page-in
        t0 = time.perf_counter()
        kp = pool.map(one, frames, chunksize=max(1, nfr // (workers * 4)))
        dt = time.perf_counter() - t0
    print(json.dumps({"value": nfr / dt, "seconds": dt, "frames": nfr, "workers": workers, "cv2": cv2.__version__,
                      "fast_keypoints_per_frame": float(np.mean(kp))}))


if __name__ == "__main__":
    main()
