"""where does a GFORB_RESIZE variant differ from the byte-gather kernel?  (diagnostic, GPU)"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gf_orb_slam_b200 import ORBextractor, synth

for (w, h, nf) in [(1241, 376, 2000), (752, 480, 1000), (200, 152, 500), (640, 480, 1000)]:
    imgs = np.stack([synth.textured_frame(w, h, seed=40 + i) for i in range(2)])
    ref = None
    for v in ("direct", "words", "words8"):
        os.environ["GFORB_RESIZE"] = v
        ex = ORBextractor(nf, 1.2, 8, max_width=w, max_height=h, max_batch=2)
        kps, desc, counts = ex.extract_batch(imgs)
        lv = [[ex.debug_level(b, l, 0).copy() for l in range(8)] for b in range(2)]
        if ref is None:
            ref = (kps.copy(), desc.copy(), counts.copy(), lv)
            continue
        for b in range(2):
            for l in range(8):
                d = np.argwhere(lv[b][l] != ref[3][b][l])
                if len(d):
                    y, x = d[0]
                    print(f"{w}x{h} {v} frame {b} level {l}: {len(d)} bytes differ, first at (y={y}, x={x}) got {lv[b][l][y, x]} want {ref[3][b][l][y, x]}; "
                          f"plane {lv[b][l].shape}, rows {sorted(set(d[:, 0]))[:6]}, cols {sorted(set(d[:, 1]))[:10]}")
        ok = np.array_equal(counts, ref[2]) and all(kps[b, :counts[b]].tobytes() == ref[0][b, :counts[b]].tobytes() and
                                                     np.array_equal(desc[b, :counts[b]], ref[1][b, :counts[b]]) for b in range(2))
        print(w, h, v, "written rows equal:", ok, "counts", counts.tolist())
