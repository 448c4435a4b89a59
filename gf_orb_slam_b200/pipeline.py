"""The per-frame tracking front end as one batched device pipeline (BASELINE.json metric: extract + match + GF-select).

For every frame t of a batch:  ORBextractor::operator() -> ORBmatcher::SearchForInitialization(frame t-1, frame t,
window 100, nnratio 0.9, checkOri)  (the map-free frame-to-frame matcher) ->
Observability::runMatrixBuilding(FRAME_INFO_MATRIX) + maxVolSelection_LazierGreedy (config C4: <=1000 matched points,
pick 100, eps 0.1).  All stages run in libgforb.so kernels on one CUDA stream; torch only owns the HBM buffers.
Frame t-1 of the first frame of a batch is the last frame of the previous batch (1-frame halo)."""
from __future__ import annotations

import os

import numpy as np
import torch

from . import synth
from .extractor import ORBextractor
from .matcher import Frames, ORBmatcher
from .observability import LAZIER_GREEDY, Observability

from .pipeline_consts import (EUROC_CAM, SIMU_XV, level_sigma2, level_sizes, sample_scale_for,  # noqa: F401
                              simu_visible_from)


class TrackingFrontEnd:
    def __init__(self, width=752, height=480, nfeatures=1000, batch=64, *, n_good=100, error_bound=0.1, window=100,
                 nnratio=0.9, device=0, cloud_seed=7, overlap=True):
        self.w, self.h, self.nf, self.B, self.k = width, height, nfeatures, batch, n_good
        self.window = window
        self.dev = torch.device("cuda", device)
        torch.cuda.set_device(self.dev)
        B, nf = batch, nfeatures
        self.ex = ORBextractor(nfeatures, 1.2, 8, max_width=width, max_height=height, max_batch=batch, device=device)
        self.matcher = ORBmatcher(nnratio, True, max_batch=batch, max_keypoints=nfeatures, max_queries=nfeatures,
                                  cand_pool=48 * nfeatures, device=device)
        cam = dict(EUROC_CAM); cam["nRows"], cam["nCols"] = height, width
        self.obs = Observability(**cam, max_batch=batch, max_points=nfeatures, device=device)
        self.obs_sel = Observability(**cam, max_batch=batch, max_points=nfeatures, device=device)   # handle used on s_sel
        if overlap:
            self.obs_sel.set_select_residency(3)    # it shares the SMs with the next step's extraction: smaller register footprint
        self.obs.set_state(SIMU_XV)
        self.obs.predictPWLSVec(0.1, 1)
        xv = self.obs.kinematic(0)["Xv"]
        self.sample_scale = sample_scale_for(nfeatures, n_good, error_bound)
        z = lambda *s, dt=torch.float32: torch.zeros(s, dtype=dt, device=self.dev)
        self.images = z(B, height, width, dt=torch.uint8)
        # keypoints / descriptors / matches are double-buffered over steps: the matcher of step t runs on its own stream
        # while the main stream already extracts step t+1 into the other buffer
        self.kps2 = [z(B + 1, nf, 7) for _ in range(2)]          # slot 0 = halo (last frame of the previous batch)
        self.desc2 = [z(B + 1, nf, 32, dt=torch.uint8) for _ in range(2)]
        self.counts2 = [z(B + 1, dt=torch.int32) for _ in range(2)]
        self.prev2 = [z(B, nf, 2) for _ in range(2)]
        self.m12_2 = [z(B, nf, dt=torch.int32) for _ in range(2)]
        self.nmatch2 = [z(B, dt=torch.int32) for _ in range(2)]
        # synthetic matched map points of every keypoint slot (tools/simu_greedy.cc recipe, fixed per slot)
        world, _, score, _ = synth.landmark_cloud(nf, cloud_seed, visible_from=simu_visible_from(width, height))
        self.world_h, self.score_h = world, score
        self.world = torch.from_numpy(np.broadcast_to(world, (B, nf, 3)).copy()).to(self.dev)
        self.score = torch.from_numpy(np.broadcast_to(score, (B, nf)).copy()).to(self.dev)
        self.valid = torch.ones((B, nf), dtype=torch.uint8, device=self.dev)
        self.octave = z(B, nf, dt=torch.int32)
        self.Xv = torch.from_numpy(np.broadcast_to(xv, (B, 13)).copy()).to(self.dev)
        self.xv_h = xv
        self.sigma2_h = level_sigma2()
        self.sigma2 = torch.from_numpy(self.sigma2_h).to(self.dev)
        # the selection runs on its own stream one step behind the extractor: its inputs and outputs are double-buffered
        self.blocks2 = [z(B, nf, 49, dt=torch.float64) for _ in range(2)]
        self.obs_score2 = [z(B, nf, dt=torch.float64) for _ in range(2)]
        self.cnt_sel2 = [z(B, dt=torch.int32) for _ in range(2)]
        self.sel_idx2 = [z(B, n_good, dt=torch.int32) for _ in range(2)]
        self.sel_ld2 = [z(B, n_good, dt=torch.float64) for _ in range(2)]
        self.sel_n2 = [z(B, dt=torch.int32) for _ in range(2)]
        self.overlap = overlap
        self.s_sel = torch.cuda.Stream(device=self.dev, priority=int(os.environ.get("GFORB_SEL_PRIORITY", "0"))) if overlap else None
        self.s_copy = torch.cuda.Stream(device=self.dev)        # device -> host results
        self.s_up = torch.cuda.Stream(device=self.dev)          # host -> device frames (its own copy engine and queue)
        self.ev_info = [torch.cuda.Event() for _ in range(2)]
        self.ev_sel = [torch.cuda.Event() for _ in range(2)]
        self.ev_h2d = [torch.cuda.Event() for _ in range(2)]
        self.ev_ext = [torch.cuda.Event() for _ in range(2)]
        self.ev_out = torch.cuda.Event()
        self.ev_d2h2 = [torch.cuda.Event() for _ in range(2)]   # host copies of buffer p's keypoints / matches are done
        self.ev_d2h_sel = [torch.cuda.Event() for _ in range(2)]
        self.t = 0                          # steps issued
        self.images2 = [self.images, z(B, height, width, dt=torch.uint8)]
        self.seeds = torch.ones(B, dtype=torch.int32, device=self.dev)
        self.F1 = [Frames(self.kps2[i][:B], self.desc2[i][:B], self.counts2[i][:B], width, height) for i in range(2)]
        self.F2 = [Frames(self.kps2[i][1:], self.desc2[i][1:], self.counts2[i][1:], width, height) for i in range(2)]
        self.kps_i32 = [k.view(torch.int32) for k in self.kps2]
        self.s_match = torch.cuda.Stream(device=self.dev, priority=int(os.environ.get("GFORB_MATCH_PRIORITY", "0"))) if overlap else None
        self.ev_extract = [torch.cuda.Event() for _ in range(2)]
        self.ev_match = [torch.cuda.Event() for _ in range(2)]
        self.launches_per_step = 0
        self.profile = False
        self.timeline = None                # list of (name, step, event) when tools/timeline.py asks for one
        self._ev = []                       # (stage, start, end) torch events recorded while self.profile
        # pinned host staging for the end-to-end path
        self.h_images = torch.zeros((B, height, width), dtype=torch.uint8).pin_memory()
        self.h_kps = torch.zeros((B, nf, 7), dtype=torch.float32).pin_memory()
        self.h_desc = torch.zeros((B, nf, 32), dtype=torch.uint8).pin_memory()
        self.h_counts = torch.zeros(B, dtype=torch.int32).pin_memory()
        self.h_m12 = torch.zeros((B, nf), dtype=torch.int32).pin_memory()
        self.h_nmatch = torch.zeros(B, dtype=torch.int32).pin_memory()
        self.h_sel_idx = torch.zeros((B, n_good), dtype=torch.int32).pin_memory()
        self.h_sel_n = torch.zeros(B, dtype=torch.int32).pin_memory()
        self._uploaded = False
        # lanes (LanedFrontEnd): the frame before this lane's first frame is the last frame of another lane
        self.halo_src = None                # TrackingFrontEnd whose slot B feeds this lane's slot 0 (None: own slot B)
        self.halo_reader = None             # TrackingFrontEnd that reads this lane's slot B as its halo
        self.ev_extracted = torch.cuda.Event()   # slot B of this lane holds the newest frame
        self.ev_halo_read = torch.cuda.Event()   # this lane has copied its halo out of halo_src's slot B

    # ---- one step = one batch of B frames through the whole front end, everything resident in HBM
    def _mark(self, name=None):
        """stage boundary: closes the open stage and opens `name` (CUDA events on the launch stream)"""
        if not self.profile:
            return
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        if self._open is not None:
            self._ev.append((self._open[0], self._open[1], e))
        self._open = (name, e) if name else None

    # the buffers of the most recently issued step (valid after sync())
    @property
    def kps(self): return self.kps2[(self.t - 1) & 1]
    @property
    def desc(self): return self.desc2[(self.t - 1) & 1]
    @property
    def counts(self): return self.counts2[(self.t - 1) & 1]
    @property
    def m12(self): return self.m12_2[(self.t - 1) & 1]
    @property
    def nmatch(self): return self.nmatch2[(self.t - 1) & 1]
    @property
    def blocks(self): return self.blocks2[(self.t - 1) & 1]
    @property
    def obs_score(self): return self.obs_score2[(self.t - 1) & 1]
    @property
    def sel_idx(self): return self.sel_idx2[(self.t - 1) & 1]
    @property
    def sel_ld(self): return self.sel_ld2[(self.t - 1) & 1]
    @property
    def sel_n(self): return self.sel_n2[(self.t - 1) & 1]

    def sync(self):
        """main stream waits for everything issued so far (selection stream, copy stream)"""
        cur = torch.cuda.current_stream()
        if self.s_sel is not None:
            for e in self.ev_sel:
                cur.wait_event(e)
        for e in self.ev_match:
            cur.wait_event(e)
        for e in self.ev_d2h2:
            cur.wait_event(e)
        for e in self.ev_d2h_sel:
            cur.wait_event(e)

    def _tl(self, name, stream):
        if self.timeline is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record(stream)
            self.timeline.append((name, self.t, e))

    def step_device(self, images: torch.Tensor | None = None):
        """extract -> match -> info on the current stream; the good-feature selection of this step on s_sel, where it
        overlaps the extraction of the next step (it is a latency-bound kernel that leaves the ALU pipe idle)"""
        B = self.B
        main = torch.cuda.current_stream()
        st = main.cuda_stream
        p = self.t & 1
        img = self.images if images is None else images
        self._open = None
        self._mark("extract")
        # halo: the frame before this batch becomes slot 0
        kps, desc, counts = self.kps2[p], self.desc2[p], self.counts2[p]
        src = self.halo_src if self.halo_src is not None else self
        if src is not self:
            main.wait_event(src.ev_extracted)
        main.wait_event(self.ev_match[p])                                   # the matcher of step t-2 is done with buffer p
        skps, sdesc, scounts = src.kps, src.desc, src.counts                # newest buffers of the halo source
        kps[0].copy_(skps[src.B]); desc[0].copy_(sdesc[src.B]); counts[0:1].copy_(scounts[src.B:src.B + 1])
        if src is not self:
            self.ev_halo_read.record(main)
        if self.halo_reader is not None and self.halo_reader is not self:
            main.wait_event(self.halo_reader.ev_halo_read)                  # the other lane has read its halo
        self._tl("extract_begin", main)
        self.ex.extract_device(img, kps[1:], desc[1:], counts[1:], stream=st, sync=False)
        self._tl("extract_end", main)
        self.ev_extracted.record(main)
        n = self.ex.last_launches()
        self._mark("match")
        # frame-to-frame matching: on its own stream when overlapping (it is a latency-bound chain of three kernels that
        # leaves the SMs almost empty; nothing downstream on the main stream consumes the matches)
        ms = self.s_match if self.s_match is not None else main
        if ms is not main:
            self.ev_extract[p].record(main)
            ms.wait_event(self.ev_extract[p])
            ms.wait_event(self.ev_d2h2[p])                                  # the host copies of step t-2's matches are done
        self._tl("match_begin", ms)
        with torch.cuda.stream(ms):
            self.prev2[p].copy_(kps[:B, :, 0:2])                            # vbPrevMatched[i] = F1.mvKeysUn[i].pt
            self.matcher.SearchForInitialization(self.F1[p], self.F2[p], self.prev2[p], self.window, matches12=self.m12_2[p],
                                                 nmatches=self.nmatch2[p], stream=ms.cuda_stream, sync=False)
            self.ev_match[p].record(ms)
        self._tl("match_end", ms)
        n += self.matcher.last_launches()
        self._mark("info")
        if self.s_sel is not None:
            main.wait_event(self.ev_sel[p])                                 # selection of step t-2 has released buffer p
        self.octave.copy_(self.kps_i32[p][1:, :, 5])
        self.cnt_sel2[p].copy_(counts[1:])
        self.obs.build_info_frame(self.Xv, self.world, self.octave, self.score, self.valid, self.cnt_sel2[p], self.sigma2,
                                  blocks=self.blocks2[p], obs_score=self.obs_score2[p], stream=st, sync=False)
        n += self.obs.last_launches()
        self._mark("select")
        if self.s_sel is not None:
            self.ev_info[p].record(main)
            self.s_sel.wait_event(self.ev_info[p])
            self.s_sel.wait_event(self.ev_d2h_sel[p])                       # results of step t-2 have left sel_*2[p]
            obs, sst = self.obs_sel, self.s_sel.cuda_stream
        else:
            main.wait_event(self.ev_d2h_sel[p])
            obs, sst = self.obs, st
        self._tl("info_end", main)
        self._tl("select_begin", self.s_sel if self.s_sel is not None else main)
        obs.select(self.blocks2[p], self.obs_score2[p], self.cnt_sel2[p], self.k, LAZIER_GREEDY, self.sample_scale, self.seeds,
                   out_idx=self.sel_idx2[p], out_logdet=self.sel_ld2[p], out_n=self.sel_n2[p], stream=sst, sync=False)
        self._tl("select_end", self.s_sel if self.s_sel is not None else main)
        n += obs.last_launches()
        if self.s_sel is not None:
            self.ev_sel[p].record(self.s_sel)
        self._mark(None)
        self.launches_per_step = n
        self.t += 1
        return n

    def stage_times(self):
        """average ms per pipeline stage over the steps recorded with self.profile (stream must be synchronised)"""
        acc = {}
        for name, e0, e1 in self._ev:
            t, c = acc.get(name, (0.0, 0))
            acc[name] = (t + e0.elapsed_time(e1), c + 1)
        self._ev = []
        return [(k, t / c) for k, (t, c) in acc.items()]

    def pyramid_bytes(self) -> int:
        return 2 * self.B * sum((w + 32) * (h + 32) for w, h in level_sizes(self.w, self.h))

    def algorithmic_bytes(self):
        """SURVEY 8(d) materialised-stage model per frame: B = 5P - P0 - P7 + 60 N, split per extractor stage."""
        px = [w * h for w, h in level_sizes(self.w, self.h)]
        P = sum(px)
        per = {"pyramid": (P - px[-1]) + (P - px[0]), "detect": P, "select": 0, "orient": 0, "blur": 2 * P,
               "describe": 60 * self.nf}
        return {"per_stage": per, "extract_frame": sum(per.values()), "traffic": {}}

    # ---- the same step through host buffers: H2D of the frames, D2H of every result a Tracking caller consumes.
    # Copies run on s_up / s_copy: the frames of step t+1 upload while step t computes (two device image buffers), results
    # download while the selection runs.
    def upload(self, q):
        """enqueue the H2D copy of self.h_images into device image buffer q on the copy stream"""
        self.s_up.wait_event(self.ev_ext[q])                                # the extraction that last read buffer q is done
        with torch.cuda.stream(self.s_up):
            self.images2[q].copy_(self.h_images, non_blocking=True)
            self.ev_h2d[q].record(self.s_up)

    def step_host(self, first=False):
        main = torch.cuda.current_stream()
        q = self.t & 1
        if first or not self._uploaded:
            self.upload(q)
        main.wait_event(self.ev_h2d[q])
        main.wait_event(self.ev_d2h2[q])                                    # step t-2's results left output buffer q
        self.step_device(self.images2[q])
        p = (self.t - 1) & 1
        self.ev_ext[q].record(main)            # (recorded after the whole main-stream part; extraction is its first stage)
        self.ev_out.record(main)
        self.upload(q ^ 1)                     # frames of the next step
        self._uploaded = True
        self.s_copy.wait_event(self.ev_out)
        self.s_copy.wait_event(self.ev_match[p])
        with torch.cuda.stream(self.s_copy):
            self.h_kps.copy_(self.kps[1:], non_blocking=True)
            self.h_desc.copy_(self.desc[1:], non_blocking=True)
            self.h_counts.copy_(self.counts[1:], non_blocking=True)
            self.h_m12.copy_(self.m12, non_blocking=True)
            self.h_nmatch.copy_(self.nmatch, non_blocking=True)
            self.ev_d2h2[p].record(self.s_copy)
        if self.s_sel is not None:
            self.s_copy.wait_event(self.ev_sel[p])
        with torch.cuda.stream(self.s_copy):
            self.h_sel_idx.copy_(self.sel_idx2[p], non_blocking=True)
            self.h_sel_n.copy_(self.sel_n2[p], non_blocking=True)
            self.ev_d2h_sel[p].record(self.s_copy)

    def h2d_bytes(self) -> int:
        return self.h_images.numel()

    def d2h_bytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in (self.h_kps, self.h_desc, self.h_counts, self.h_m12, self.h_nmatch,
                                                           self.h_sel_idx, self.h_sel_n))


class LanedFrontEnd:
    """The same front end as `lanes` TrackingFrontEnd pipelines of batch/lanes frames each, one CUDA stream per lane:
    while one lane runs its latency-bound tail (quota / retainBest emulation, candidate resolution) the other lane's
    ALU-bound kernels (pyramid, FAST, blur) fill the SMs.  Lane j takes frames [j n, (j+1) n) of every batch; its halo
    frame is the last frame of lane j-1 (of the previous batch for lane 0), handed over with events, so the results are
    those of one TrackingFrontEnd over the whole batch."""

    def __init__(self, width=752, height=480, nfeatures=1000, batch=64, lanes=2, **kw):
        assert batch % lanes == 0
        self.B, self.L, self.n = batch, lanes, batch // lanes
        self.w, self.h, self.nf = width, height, nfeatures
        dev = kw.get("device", 0)
        self.dev = torch.device("cuda", dev)
        torch.cuda.set_device(self.dev)
        self.lane = [TrackingFrontEnd(width, height, nfeatures, self.n, **kw) for _ in range(lanes)]
        # the lane (extraction) streams are high priority: their short ALU-bound CTAs take freed SM slots before the
        # long-lived CTAs of the selection / matching streams do (3.87 -> 3.68 ms per step on B200)
        self.streams = [torch.cuda.Stream(device=self.dev, priority=int(os.environ.get("GFORB_MAIN_PRIORITY", "-3"))) for _ in range(lanes)]
        for j, fe in enumerate(self.lane):
            fe.halo_src = self.lane[(j - 1) % lanes]
            fe.halo_reader = self.lane[(j + 1) % lanes]
        self.ev_fork = torch.cuda.Event()
        self.ev_join = [torch.cuda.Event() for _ in range(lanes)]
        self.k, self.sample_scale = self.lane[0].k, self.lane[0].sample_scale
        self.launches_per_step = 0
        self._forked = False

    # bench.py / smoke interface ------------------------------------------------------------------------------------
    @property
    def profile(self): return self.lane[0].profile
    @profile.setter
    def profile(self, v):
        for fe in self.lane: fe.profile = v

    def set_extract_profiling(self, on):
        for fe in self.lane: fe.ex.set_profiling(on)

    def set_overlap(self, on):
        """False: the good-feature selection of every lane runs on the lane's main stream"""
        for fe in self.lane:
            if not hasattr(fe, "_s_sel_saved"): fe._s_sel_saved = fe.s_sel
            fe.s_sel = fe._s_sel_saved if on else None

    def load_host_frames(self, host: np.ndarray):
        for j, fe in enumerate(self.lane):
            fe.h_images.copy_(torch.from_numpy(host[j * self.n:(j + 1) * self.n]))
            fe.images.copy_(fe.h_images)

    def _fork(self):
        """lane streams start after everything already queued on the caller's stream (once per burst of steps)"""
        if not self._forked:
            self.ev_fork.record(torch.cuda.current_stream())
            for s in self.streams: s.wait_event(self.ev_fork)
            self._forked = True

    def step_device(self):
        self._fork()
        n = 0
        for fe, s in zip(self.lane, self.streams):
            with torch.cuda.stream(s):
                n += fe.step_device()
        self.launches_per_step = n
        return n

    def step_host(self):
        self._fork()
        for fe, s in zip(self.lane, self.streams):
            with torch.cuda.stream(s):
                fe.step_host()

    def sync(self):
        """the caller's stream waits for every lane (and the lanes' selection / copy streams)"""
        cur = torch.cuda.current_stream()
        for j, (fe, s) in enumerate(zip(self.lane, self.streams)):
            with torch.cuda.stream(s):
                fe.sync()
                self.ev_join[j].record(s)
            cur.wait_event(self.ev_join[j])
        self._forked = False

    def stage_times(self):
        return _sum_stage_times([fe.stage_times() for fe in self.lane])

    def extract_stage_times(self):
        return _sum_stage_times([fe.ex.stage_times() for fe in self.lane])

    def pyramid_bytes(self): return sum(fe.pyramid_bytes() for fe in self.lane)
    def algorithmic_bytes(self): return self.lane[0].algorithmic_bytes()
    def h2d_bytes(self): return sum(fe.h2d_bytes() for fe in self.lane)
    def d2h_bytes(self): return sum(fe.d2h_bytes() for fe in self.lane)


def _sum_stage_times(per_lane):
    """[(stage, ms)] per lane -> [(stage, ms summed over the lanes)]: with L lanes of B/L frames the sum is the time the
    stage's kernels take for the whole batch"""
    acc = {}
    for st in per_lane:
        for k, t in st:
            acc[k] = acc.get(k, 0.0) + t
    return list(acc.items())
