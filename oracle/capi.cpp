// TEST INFRASTRUCTURE ONLY (see oracle.hpp).  extern "C" surface so tests/ and bench.py's cpu_baseline leg can
// drive the oracle through ctypes.  Not part of the product ABI (that is include/gforb.h).
#include "oracle.hpp"
#include <cmath>
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <thread>

using namespace oracle;

// rand() of a seeded selection is a thread-local random_r (the same glibc TYPE_3 generator): callers may run selections on
// several threads (tests, front_end_batch) without racing on the process-global rand() state.
#include <stdlib.h>
namespace {
thread_local struct random_data tl_rd;
thread_local char tl_state[128];
thread_local bool tl_init = false;
int tl_rand() { int32_t r; random_r(&tl_rd, &r); return (int)r; }
void tl_srand(unsigned seed) {
    if (!tl_init) { memset(&tl_rd, 0, sizeof(tl_rd)); initstate_r(seed, tl_state, sizeof(tl_state), &tl_rd); tl_init = true; }
    srandom_r(seed, &tl_rd);
}
}  // namespace

extern "C" {

// ------------------------------------------------------------------ primitives
int orc_cvround(double v) { return cvRound(v); }
void orc_set_fast_scalar(int on) { oracle::g_fast_scalar_only = on != 0; }
void orc_fastatan2(const float* y, const float* x, float* out, int n) { for (int i = 0; i < n; i++) out[i] = fastAtan2(y[i], x[i]); }
void orc_resize(const uint8_t* src, int sw, int sh, int sstep, uint8_t* dst, int dw, int dh, int dstep) { resizeLinear8u(src, sw, sh, sstep, dst, dw, dh, dstep); }
void orc_border101(const uint8_t* src, int w, int h, int sstep, uint8_t* dst, int dstep, int b) { copyMakeBorder101(src, w, h, sstep, dst, dstep, b); }
void orc_blur(const uint8_t* src, int w, int h, int sstep, uint8_t* dst, int dstep, int variant) { gaussianBlur7x7(src, w, h, sstep, dst, dstep, variant); }
int orc_fast(const uint8_t* img, int w, int h, int step, int th, int nonmax, KeyPoint* out, int cap) {
    std::vector<KeyPoint> k;
    FAST9_16(img, w, h, step, th, nonmax != 0, k);
    int n = (int)std::min<size_t>(k.size(), cap);
    memcpy(out, k.data(), sizeof(KeyPoint) * n);
    return (int)k.size();
}
// responses in; out_idx = indices (into the input) of the surviving keypoints in their output order
int orc_retain_best(const float* resp, int n_in, int n_keep, int variant, int* out_idx) {
    std::vector<KeyPoint> k(n_in);
    for (int i = 0; i < n_in; i++) { k[i] = KeyPoint{0, 0, 0, 0, resp[i], 0, i}; }
    retainBest(k, n_keep, variant);
    for (size_t i = 0; i < k.size(); i++) out_idx[i] = k[i].class_id;
    return (int)k.size();
}
void orc_sincosf(const float* x, float* s, float* c, int n) { for (int i = 0; i < n; i++) { s[i] = sinf(x[i]); c[i] = cosf(x[i]); } }
void orc_rand_stream(unsigned seed, int n, int* out) { srand(seed); for (int i = 0; i < n; i++) out[i] = rand(); }

// ------------------------------------------------------------------ extractor
void* orc_extractor_create(int nfeatures, float scaleFactor, int nlevels, int scoreType, int fastTh, int minFastTh,
                           int blurVariant, int retainVariant) {
    return new ORBextractor(nfeatures, scaleFactor, nlevels, scoreType, fastTh, minFastTh, blurVariant, retainVariant);
}
void orc_extractor_destroy(void* h) { delete (ORBextractor*)h; }
int orc_extractor_run(void* h, const uint8_t* img, int w, int ht, int step, KeyPoint* kps, uint8_t* desc, int cap) {
    ORBextractor* e = (ORBextractor*)h;
    std::vector<KeyPoint> k; std::vector<uint8_t> d;
    int rc = (*e)(img, w, ht, step, k, d);
    if (rc < 0) return rc;
    int n = (int)std::min<size_t>(k.size(), cap);
    if (kps) memcpy(kps, k.data(), sizeof(KeyPoint) * n);
    if (desc) memcpy(desc, d.data(), (size_t)32 * n);
    return (int)k.size();
}
// many frames, `threads` worker threads, one extractor instance per thread (the reference extractor is single-threaded)
int orc_extractor_run_batch(int nfeatures, float scaleFactor, int nlevels, int fastTh, int minFastTh, int blurVariant,
                            int retainVariant, const uint8_t* imgs, int w, int ht, int nframes, int threads,
                            KeyPoint* kps, uint8_t* desc, int* counts, int cap) {
    if (threads < 1) threads = 1;
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++)
        pool.emplace_back([=]() {
            ORBextractor e(nfeatures, scaleFactor, nlevels, FAST_SCORE, fastTh, minFastTh, blurVariant, retainVariant);
            std::vector<KeyPoint> k; std::vector<uint8_t> d;
            for (int f = t; f < nframes; f += threads) {
                e(imgs + (size_t)f * w * ht, w, ht, w, k, d);
                int n = (int)std::min<size_t>(k.size(), cap);
                counts[f] = n;
                if (kps) memcpy(kps + (size_t)f * cap, k.data(), sizeof(KeyPoint) * n);
                if (desc) memcpy(desc + (size_t)f * cap * 32, d.data(), (size_t)32 * n);
            }
        });
    for (auto& th : pool) th.join();
    return 0;
}
int orc_extractor_pyramid(void* h, const uint8_t* img, int w, int ht, int step) {
    ((ORBextractor*)h)->ComputePyramid(img, w, ht, step); return 0;
}
void orc_extractor_level_info(void* h, int level, int* w, int* ht, int* step, int* nfeat) {
    ORBextractor* e = (ORBextractor*)h;
    *w = e->mvImagePyramid[level].w; *ht = e->mvImagePyramid[level].h; *step = e->mvImagePyramid[level].step;
    *nfeat = e->mnFeaturesPerLevel[level];
}
void orc_extractor_level_copy(void* h, int level, uint8_t* dst) {
    ORBextractor* e = (ORBextractor*)h;
    memcpy(dst, e->mvImagePyramid[level].buf.data(), e->mvImagePyramid[level].buf.size());
}
void orc_extractor_tables(void* h, float* scale, float* invScale, int* umax16) {
    ORBextractor* e = (ORBextractor*)h;
    for (int i = 0; i < e->nlevels; i++) { scale[i] = e->mvScaleFactor[i]; invScale[i] = e->mvInvScaleFactor[i]; }
    for (int i = 0; i < 16; i++) umax16[i] = e->umax[i];
}
// per-level keypoints (level coordinates, angle filled) of the last run
int orc_extractor_level_keypoints(void* h, int level, KeyPoint* out, int cap) {
    ORBextractor* e = (ORBextractor*)h;
    if (level >= (int)e->lastAllKeypoints.size()) return 0;
    auto& v = e->lastAllKeypoints[level];
    int n = (int)std::min<size_t>(v.size(), cap);
    memcpy(out, v.data(), sizeof(KeyPoint) * n);
    return (int)v.size();
}

// ------------------------------------------------------------------ frame / matcher
void* orc_frame_create(const KeyPoint* kps, const uint8_t* desc, int n, int imgW, int imgH, int nlevels, float scaleFactor) {
    Frame* f = new Frame();
    f->init(kps, desc, n, imgW, imgH, nlevels, scaleFactor);
    return f;
}
void orc_frame_destroy(void* f) { delete (Frame*)f; }
void orc_frame_set_camera(void* fp, float fx, float fy, float cx, float cy, const float* Tcw) {
    Frame* f = (Frame*)fp; f->fx = fx; f->fy = fy; f->cx = cx; f->cy = cy; if (Tcw) memcpy(f->mTcw, Tcw, 64);
}
void orc_frame_set_mappoints(void* fp, const int* mp, const uint8_t* outlier) {
    Frame* f = (Frame*)fp;
    for (int i = 0; i < f->N; i++) { f->mvpMapPoints[i] = mp[i]; if (outlier) f->mvbOutlier[i] = outlier[i]; }
}
void orc_frame_get_matches(void* fp, int* mp, int* score) {
    Frame* f = (Frame*)fp;
    for (int i = 0; i < f->N; i++) { mp[i] = f->mvpMapPoints[i]; score[i] = f->mvpMatchScore[i]; }
}
int orc_frame_features_in_area(void* fp, float x, float y, float r, int minL, int maxL, int* out, int cap) {
    auto v = ((Frame*)fp)->GetFeaturesInArea(x, y, r, minL, maxL);
    for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = (int)v[i];
    return (int)v.size();
}
int orc_descriptor_distance(const uint8_t* a, const uint8_t* b) { return ORBmatcher::DescriptorDistance(a, b); }

// map-point SoA: flags bit0 = mbTrackInView, bit1 = isBad
int orc_search_by_projection(void* fp, int M, const float* u, const float* v, const int* level, const float* viewCos,
                             const uint8_t* flags, const uint8_t* desc, float th, float nnratio) {
    std::vector<MapPointView> mps(M);
    for (int i = 0; i < M; i++) {
        mps[i].mTrackProjX = u[i]; mps[i].mTrackProjY = v[i]; mps[i].mnTrackScaleLevel = level[i];
        mps[i].mTrackViewCos = viewCos[i]; mps[i].mbTrackInView = flags[i] & 1; mps[i].isBad = (flags[i] >> 1) & 1;
        memcpy(mps[i].desc, desc + (size_t)32 * i, 32);
    }
    ORBmatcher m(nnratio, true);
    return m.SearchByProjection(*(Frame*)fp, mps, th);
}
// SearchByProjection_GoodFeature(F, MPs, th_good, th_rest): flags bit2 = goodAtFrameId == F.mnId
int orc_search_by_projection_good(void* fp, int M, const float* u, const float* v, const int* level, const float* viewCos,
                                  const uint8_t* flags, const uint8_t* desc, float th_good, float th_rest, float nnratio) {
    std::vector<MapPointView> mps(M);
    std::vector<uint8_t> good(M);
    for (int i = 0; i < M; i++) {
        mps[i].mTrackProjX = u[i]; mps[i].mTrackProjY = v[i]; mps[i].mnTrackScaleLevel = level[i];
        mps[i].mTrackViewCos = viewCos[i]; mps[i].mbTrackInView = flags[i] & 1; mps[i].isBad = (flags[i] >> 1) & 1;
        good[i] = (flags[i] >> 2) & 1;
        memcpy(mps[i].desc, desc + (size_t)32 * i, 32);
    }
    ORBmatcher m(nnratio, true);
    return m.SearchByProjection_GoodFeature(*(Frame*)fp, mps, good.data(), th_good, th_rest);
}
int orc_search_by_projection_frames_good(void* cur, void* last, const float* lastWorldPos, const uint8_t* good, float th_good,
                                         float th_rest, float nnratio, int checkOri) {
    ORBmatcher m(nnratio, checkOri != 0);
    return m.SearchByProjection_GoodFeature(*(Frame*)cur, *(const Frame*)last, lastWorldPos, good, th_good, th_rest);
}
int orc_search_by_projection_frames(void* cur, void* last, const float* lastWorldPos, float th, float nnratio, int checkOri) {
    ORBmatcher m(nnratio, checkOri != 0);
    return m.SearchByProjection(*(Frame*)cur, *(const Frame*)last, lastWorldPos, th);
}
int orc_search_for_initialization(void* f1, void* f2, float* prevMatched, int* matches12, int windowSize, float nnratio, int checkOri) {
    Frame* F1 = (Frame*)f1;
    ORBmatcher m(nnratio, checkOri != 0);
    std::vector<float> pm(prevMatched, prevMatched + 2 * (size_t)F1->N);
    std::vector<int> m12;
    int n = m.SearchForInitialization(*F1, *(Frame*)f2, pm, m12, windowSize);
    memcpy(prevMatched, pm.data(), sizeof(float) * pm.size());
    memcpy(matches12, m12.data(), sizeof(int) * m12.size());
    return n;
}
int orc_search_by_projection_window(void* f1, void* f2, const float* f1World, int windowSize, int* matches2, float nnratio) {
    ORBmatcher m(nnratio, true);
    std::vector<int> m2;
    int n = m.SearchByProjection(*(const Frame*)f1, *(Frame*)f2, f1World, windowSize, m2);
    memcpy(matches2, m2.data(), sizeof(int) * m2.size());
    return n;
}
void orc_extractor_set_distribution(void* e, int mode) { ((ORBextractor*)e)->distribution = mode; }
float orc_ic_angle_at(const uint8_t* img, int step, float x, float y) { return ic_angle_at(img, step, x, y); }
void orc_describe_at(const uint8_t* img, int step, float x, float y, float angleDeg, uint8_t* desc) { describe_at(img, step, x, y, angleDeg, desc); }
int orc_search_by_bow(void* kf, const int* kfNode, void* f, const int* fNode, int* matches, float nnratio, int checkOri) {
    ORBmatcher m(nnratio, checkOri != 0);
    std::vector<int> mm;
    int n = m.SearchByBoW(*(const Frame*)kf, kfNode, *(const Frame*)f, fNode, mm);
    memcpy(matches, mm.data(), sizeof(int) * mm.size());
    return n;
}
int orc_search_by_projection_keyframe(void* cur, void* kf, const uint8_t* kfSkip, const float* kfWorld, const float* kfMinDist,
                                      const uint8_t* kfMpDesc, float th, int ORBdist, float nnratio, int checkOri) {
    ORBmatcher m(nnratio, checkOri != 0);
    return m.SearchByProjection(*(Frame*)cur, *(const Frame*)kf, kfSkip, kfWorld, kfMinDist, kfMpDesc, th, ORBdist);
}
// ---- mapping / loop-closing overloads (orb_matcher_kf.cpp).  Map-point geometry as SoA arrays indexed by id.
static std::vector<MapPointGeom> geom_of(int nIds, const float* world, const float* normal, const float* minDist, const float* maxDist,
                                         const uint8_t* desc) {
    std::vector<MapPointGeom> g(nIds);
    for (int i = 0; i < nIds; i++) {
        memcpy(g[i].world, world + 3 * (size_t)i, 12);
        if (normal) memcpy(g[i].normal, normal + 3 * (size_t)i, 12); else g[i].normal[0] = g[i].normal[1] = g[i].normal[2] = 0;
        g[i].minDist = minDist[i]; g[i].maxDist = maxDist[i];
        memcpy(g[i].desc, desc + 32 * (size_t)i, 32);
    }
    return g;
}
void orc_cv_rx_plus_t(const float* R9, const float* x3, const float* t3, float* y3) { cvGemmRxPlusT(R9, 3, x3, t3, 1, y3); }
void orc_cv_neg_rt_t(const float* R9, const float* t3, float* Ow3) { cvNegRtTimesT(R9, 3, t3, 1, Ow3); }
void orc_sim3_decompose(const float* Scw16, float* R9, float* t3, float* Ow3) {
    const Sim3Rt S = decomposeSim3(Scw16);
    memcpy(R9, S.R, 36); memcpy(t3, S.t, 12); memcpy(Ow3, S.Ow, 12);
}
int orc_search_by_bow_kf(void* kf1, const int* node1, void* kf2, const int* node2, int* matches12, float nnratio, int checkOri) {
    ORBmatcher m(nnratio, checkOri != 0);
    std::vector<int> mm;
    const int n = m.SearchByBoW_KeyFrames(*(const Frame*)kf1, node1, *(const Frame*)kf2, node2, mm);
    memcpy(matches12, mm.data(), sizeof(int) * mm.size());
    return n;
}
int orc_search_for_triangulation(void* kf1, const int* node1, void* kf2, const int* node2, const float* F12, int* matches12, int checkOri) {
    ORBmatcher m(0.6f, checkOri != 0);
    std::vector<int> mm;
    const int n = m.SearchForTriangulation(*(const Frame*)kf1, node1, *(const Frame*)kf2, node2, F12, mm);
    memcpy(matches12, mm.data(), sizeof(int) * mm.size());
    return n;
}
// Scw == nullptr: Fuse(pKF, vpMapPoints, th) with the pose in the frame (orc_frame_set_camera); else Fuse(pKF, Scw, vpPoints, th)
int orc_fuse(void* kf, const float* Scw, const int* points, int nPoints, int nIds, const float* world, const float* normal,
             const float* minDist, const float* maxDist, const uint8_t* desc, float th, int* kfSlot, uint8_t* bad, int* idxInKF,
             int* bestIdx, int* bestDist) {
    const Frame& KF = *(const Frame*)kf;
    const std::vector<MapPointGeom> g = geom_of(nIds, world, normal, minDist, maxDist, desc);
    std::vector<int> pts(points, points + nPoints), slot(kfSlot, kfSlot + KF.N), inKF(idxInKF, idxInKF + nIds), bi, bd;
    std::vector<uint8_t> b(bad, bad + nIds);
    ORBmatcher m;
    const int n = Scw ? m.Fuse(KF, Scw, pts, g.data(), th, slot, b, inKF, bi, bd) : m.Fuse(KF, pts, g.data(), th, slot, b, inKF, bi, bd);
    memcpy(kfSlot, slot.data(), sizeof(int) * slot.size()); memcpy(bad, b.data(), b.size()); memcpy(idxInKF, inKF.data(), sizeof(int) * inKF.size());
    memcpy(bestIdx, bi.data(), sizeof(int) * bi.size()); memcpy(bestDist, bd.data(), sizeof(int) * bd.size());
    return n;
}
int orc_search_by_projection_sim3(void* kf, const float* Scw, const int* points, int nPoints, int nIds, const float* world,
                                  const float* normal, const float* minDist, const float* maxDist, const uint8_t* desc,
                                  const uint8_t* bad, int* matched, int th) {
    const Frame& KF = *(const Frame*)kf;
    const std::vector<MapPointGeom> g = geom_of(nIds, world, normal, minDist, maxDist, desc);
    std::vector<int> pts(points, points + nPoints), mt(matched, matched + KF.N);
    ORBmatcher m;
    const int n = m.SearchByProjection(KF, Scw, pts, g.data(), bad, mt, th);
    memcpy(matched, mt.data(), sizeof(int) * mt.size());
    return n;
}
int orc_search_by_sim3(void* kf1, void* kf2, int* matches12, const int* idx2InKF2, int nIds, const float* world, const float* minDist,
                       const float* maxDist, const uint8_t* desc, const uint8_t* bad, float s12, const float* R12, const float* t12,
                       float th) {
    const Frame& KF1 = *(const Frame*)kf1;
    const std::vector<MapPointGeom> g = geom_of(nIds, world, nullptr, minDist, maxDist, desc);
    std::vector<int> m12(matches12, matches12 + KF1.N);
    ORBmatcher m;
    const int n = m.SearchBySim3(KF1, *(const Frame*)kf2, m12, idx2InKF2, g.data(), bad, s12, R12, t12, th);
    memcpy(matches12, m12.data(), sizeof(int) * m12.size());
    return n;
}
int orc_window_search(void* f1, void* f2, int windowSize, int* matches2, int minLevel, int maxLevel, float nnratio, int checkOri) {
    ORBmatcher m(nnratio, checkOri != 0);
    std::vector<int> m2;
    int n = m.WindowSearch(*(Frame*)f1, *(Frame*)f2, windowSize, m2, minLevel, maxLevel);
    memcpy(matches2, m2.data(), sizeof(int) * m2.size());
    return n;
}

// ------------------------------------------------------------------ observability
void* orc_obs_create_f(double f, int nRows, int nCols, double Cx, double Cy, double k1, double k2, double dx, double dy) {
    return new Observability(f, nRows, nCols, Cx, Cy, k1, k2, dx, dy);
}
void* orc_obs_create_fufv(double fu, double fv, int nRows, int nCols, double Cx, double Cy, double k1, double k2) {
    return new Observability(fu, fv, nRows, nCols, Cx, Cy, k1, k2);
}
void orc_obs_destroy(void* o) { delete (Observability*)o; }
void orc_obs_set_xv(void* o, const double* Xv) { memcpy(((Observability*)o)->Xv, Xv, 13 * sizeof(double)); }
void orc_obs_get_xv(void* o, double* Xv) { memcpy(Xv, ((Observability*)o)->Xv, 13 * sizeof(double)); }
void orc_obs_update(void* o, double t0, const float* Tcw_prev, double t1, const float* Twc_cur) { ((Observability*)o)->updatePWLSVec(t0, Tcw_prev, t1, Twc_cur); }
void orc_obs_predict(void* o, double dt, int nseg) { ((Observability*)o)->predictPWLSVec(dt, nseg); }
void orc_obs_get_kine(void* o, int seg, double* Xv, double* F_Q, double* F_Omg, double* F, double* F_Q_inSeg,
                      double* F_Omg_inSeg, float* Tcw) {
    KineStruct& k = ((Observability*)o)->kinematic[seg];
    memcpy(Xv, k.Xv, sizeof(k.Xv)); memcpy(F_Q, k.F_Q, sizeof(k.F_Q)); memcpy(F_Omg, k.F_Omg, sizeof(k.F_Omg));
    memcpy(F, k.F, sizeof(k.F)); memcpy(F_Q_inSeg, k.F_Q_inSeg, sizeof(k.F_Q_inSeg));
    memcpy(F_Omg_inSeg, k.F_Omg_inSeg, sizeof(k.F_Omg_inSeg)); memcpy(Tcw, k.Tcw, sizeof(k.Tcw));
}
void orc_obs_H_complete(void* o, const double* Xv, const double* yi, const double* zi, double* H13, double* H47) {
    double dh[6]; float ru, rv;
    ((Observability*)o)->compute_H_subblock_complete(Xv, yi, zi, H13, H47, dh, ru, rv);
}
int orc_obs_H_simplified(void* o, const double* Xv, const double* yi, double* H13, double* H47, float* uv) {
    double dh[6];
    return ((Observability*)o)->compute_H_subblock_simplied(Xv, yi, H13, H47, dh, false, uv[0], uv[1]);
}
void orc_obs_build_frame(void* o, int seg, const float* worldPos, const int* octave, const int* matchScore,
                         const uint8_t* valid, int n, const float* levelSigma2, double* blocks, double* obsScore, double* H_meas) {
    Observability* ob = (Observability*)o; ob->mKineIdx = seg;
    ob->batchInfoMat_Frame(worldPos, octave, matchScore, valid, n, levelSigma2, blocks, obsScore, H_meas);
}
void orc_obs_build_map(void* o, int seg, const float* worldPos, const uint8_t* trackInView, int n, int check_viz,
                       int minX, int maxX, int minY, int maxY, double* blocks, double* obsScore, float* uv, double* H_meas) {
    Observability* ob = (Observability*)o; ob->mKineIdx = seg;
    ob->batchInfoMat_Map(worldPos, trackInView, n, check_viz != 0, minX, maxX, minY, maxY, blocks, obsScore, uv, H_meas);
}
double orc_logdet(const double* M, int n) { return logDet(M, n); }
// mtd: 1 baseline greedy, 2 lazier greedy selection, 3 automatic (deletion if 2k>n), 5 deletion.
// seed >= 0 -> srand(seed) first (glibc default stream == seed 1).  Returns number selected; out_idx = pool indices
// in selection order (deletion: ascending pool order); out_logdet[i] = log-det after pick i (selection modes).
int orc_obs_select(const double* blocks, const double* obsScore, int n, int k, int mtd, double sampleScale, long seed,
                   int* out_idx, double* out_logdet) {
    Observability ob(1.0, 1.0, 480, 640, 320, 240, 0, 0);
    ob.lmkSelectPool.resize(n);
    for (int i = 0; i < n; i++) {
        GoodPoint& g = ob.lmkSelectPool[i];
        g.idx = i; g.obs_score = obsScore ? obsScore[i] : 1.0; g.selected = false; g.upper_bound = 0;
        memcpy(g.obs_block, blocks + (size_t)49 * i, sizeof(g.obs_block));
    }
    if (seed >= 0) { tl_srand((unsigned)seed); ob.rand_fn = &tl_rand; }      // seed < 0: the process-global rand() as it stands
    std::vector<GoodPoint> sub;
    if (mtd == 1) ob.maxVolSelection_BaselineGreedy(&sub, k);
    else if (mtd == 2) ob.maxVolSelection_LazierGreedy(0, n, &sub, k, sampleScale, 1e30);
    else if (mtd == 3) ob.maxVolAutomatic_LazierGreedy(0, n, &sub, k, sampleScale, 1e30);
    else if (mtd == 5) ob.maxVolDeletion_LazierGreedy(0, n, &sub, k, sampleScale, 1e30);
    else return -1;
    for (size_t i = 0; i < sub.size(); i++) { out_idx[i] = (int)sub[i].idx; if (out_logdet) out_logdet[i] = sub[i].upper_bound; }
    return (int)sub.size();
}

// Selection stage of the local-map setSelction_Number with greedy_mtd 3 (src/Observability.cc:1085-1163): pool of n entries
// (all with ObsScore >= 0), Greedy_Paral_TrigSz 1000, Greedy_Paral_RelFac 1.2, random_sample_scale 6.0, numThreads =
// mNumThreads.  The reference runs the ranges on std::threads sharing the global rand() (timing dependent); this
// restatement -- like the product -- gives range i the stream seed + 1 + i and the final pass the stream `seed`.
int orc_obs_select_map_mtd3(const double* blocks, const double* obsScore, int n, int k, long seed, int numThreads, int* out_idx) {
    Observability ob(1.0, 1.0, 480, 640, 320, 240, 0, 0);
    ob.lmkSelectPool.resize(n);
    for (int i = 0; i < n; i++) {
        GoodPoint& g = ob.lmkSelectPool[i];
        g.idx = i; g.obs_score = obsScore ? obsScore[i] : 1.0; g.selected = false; g.upper_bound = 0;
        memcpy(g.obs_block, blocks + (size_t)49 * i, sizeof(g.obs_block));
    }
    ob.rand_fn = &tl_rand;
    const double random_sample_scale = 6.0;
    const size_t num_good_inlier = (size_t)k;
    std::vector<GoodPoint> mpVec;
    size_t threadNeeded = 1;
    if (ob.lmkSelectPool.size() - 1.2 * float(num_good_inlier) <= 10 || ob.lmkSelectPool.size() < 2 * 1000) threadNeeded = 1;
    else {
        threadNeeded = static_cast<size_t>(round(float(ob.lmkSelectPool.size()) / float(1000)));
        if (threadNeeded > (size_t)numThreads) threadNeeded = (size_t)numThreads;
    }
    if (threadNeeded == 1) {
        tl_srand((unsigned)seed);
        ob.maxVolAutomatic_LazierGreedy(0, ob.lmkSelectPool.size(), &mpVec, num_good_inlier, random_sample_scale, 1e30);
    } else {
        size_t numSubsetParal = static_cast<size_t>(ceil(float(num_good_inlier) / float(threadNeeded) * 1.2));
        size_t numPoolParal = static_cast<size_t>(ceil(float(ob.lmkSelectPool.size()) / float(threadNeeded)));
        std::vector<std::vector<GoodPoint>> parVec(threadNeeded);
        for (size_t i = 0; i < threadNeeded; i++) {
            tl_srand((unsigned)(seed + 1 + (long)i));
            const size_t st = i * numPoolParal, ed = (i + 1 < threadNeeded) ? (i + 1) * numPoolParal : ob.lmkSelectPool.size();
            ob.maxVolAutomatic_LazierGreedy(st, ed, &parVec[i], numSubsetParal, random_sample_scale, 1e30);
        }
        ob.lmkSelectPool.clear();
        for (size_t i = 0; i < threadNeeded; i++)
            for (size_t j = 0; j < parVec[i].size(); ++j) { parVec[i][j].selected = false; ob.lmkSelectPool.push_back(parVec[i][j]); }
        tl_srand((unsigned)seed);
        ob.maxVolAutomatic_LazierGreedy(0, ob.lmkSelectPool.size(), &mpVec, num_good_inlier, random_sample_scale, 1e30);
    }
    for (size_t i = 0; i < mpVec.size(); i++) out_idx[i] = (int)mpVec[i].idx;
    return (int)mpVec.size();
}

int orc_pose_optimization(float* Tcw, int N, const float* kpXY, const int* octave, const int* mapPoint, const float* world,
                          const float* invLevelSigma2, float fx, float fy, float cx, float cy, uint8_t* outlier) {
    return PoseOptimization(Tcw, N, kpXY, octave, mapPoint, world, invLevelSigma2, fx, fy, cx, cy, outlier);
}

// Frame::isInFrustum for M map points (src/Frame.cc:166-227); flags[i] bit0 <- mbTrackInView
void orc_is_in_frustum(const float* Rcw, const float* tcw, const float* Ow, float fx, float fy, float cx, float cy, int minX,
                       int maxX, int minY, int maxY, int nlevels, float scaleFactor, int M, const float* world,
                       const float* normal, const float* minDist, const float* maxDist, float cosLimit, float* u, float* v,
                       int* level, float* viewCos, uint8_t* flags) {
    std::vector<float> sf(nlevels, 1.0f);
    for (int i = 1; i < nlevels; i++) sf[i] = sf[i - 1] * scaleFactor;
    for (int i = 0; i < M; i++) {
        const bool in = isInFrustum(Rcw, tcw, Ow, fx, fy, cx, cy, minX, maxX, minY, maxY, sf, world + 3 * i, normal + 3 * i,
                                    minDist[i], maxDist[i], cosLimit, u[i], v[i], level[i], viewCos[i]);
        flags[i] = (uint8_t)((flags[i] & ~1) | (in ? 1 : 0));
    }
}

// MapPoint::ComputeDistinctiveDescriptors for n_points map points (concatenated observation descriptors + offsets)
void orc_distinctive_descriptors(int n_points, const uint8_t* descriptors, const int* offsets, int* best_idx) {
    for (int i = 0; i < n_points; i++)
        best_idx[i] = distinctiveDescriptor(descriptors + (size_t)32 * offsets[i], offsets[i + 1] - offsets[i]);
}

// Frame::UndistortKeyPoints (src/Frame.cc:389-423) on n x 2 positions; dist[0] == 0 -> copy (the early return)
void orc_undistort_keypoints(const float* src, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist,
                             float* dst) {
    if (ndist <= 0 || dist[0] == 0.0f) { for (int i = 0; i < 2 * n; i++) dst[i] = src[i]; return; }
    undistortPoints(src, n, fx, fy, cx, cy, dist, ndist, dst);
}

// Observability::runActiveMapMatching (src/Observability.cc:1249-1524).  Map-point SoA as in orc_search_by_projection;
// in_pool[i] = (updateAtFrameId == frame id); cur_mat (49) in: mBaseInfoMat, out: accumulated matrix; the frame's
// mvpMapPoints / mvpMatchScore are updated in place; left_idx receives mLeftMapPoints, *left_n its size.
int orc_active_match(void* fp, int M, const float* u, const float* v, const int* level, const float* viewCos,
                     const uint8_t* flags, const uint8_t* desc, const uint8_t* in_pool, const double* obsMat,
                     const double* H_meas, double* cur_mat, float th, float nnratio, int num_to_match, long seed,
                     int* left_idx, int* left_n) {
    std::vector<MapPointView> mps(M);
    for (int i = 0; i < M; i++) {
        mps[i].mTrackProjX = u[i]; mps[i].mTrackProjY = v[i]; mps[i].mnTrackScaleLevel = level[i];
        mps[i].mTrackViewCos = viewCos[i]; mps[i].mbTrackInView = flags[i] & 1; mps[i].isBad = (flags[i] >> 1) & 1;
        memcpy(mps[i].desc, desc + (size_t)32 * i, 32);
    }
    ORBmatcher m(nnratio, true);
    if (seed >= 0) srand((unsigned)seed);
    std::vector<int> left;
    int n = runActiveMapMatching(*(Frame*)fp, mps, in_pool, obsMat, H_meas, cur_mat, th, m, num_to_match, left);
    for (size_t i = 0; i < left.size(); i++) left_idx[i] = left[i];
    *left_n = (int)left.size();
    return n;
}

}  // extern "C"

// ------------------------------------------------------------------ whole front end (cpu_baseline / --impl reference)
// extract -> SearchForInitialization(prev, cur) -> batchInfoMat_Frame + maxVolSelection_LazierGreedy for every frame,
// frames split into contiguous chunks over `threads` std::threads (one extractor per thread, 1-frame halo per chunk:
// the reference extractor/matcher are single-threaded per frame).  rand() is replaced by a thread-local random_r
// (same glibc TYPE_3 generator, seed 1 per frame) so threads neither race nor contend on glibc's lock.

extern "C" int orc_front_end_batch(const uint8_t* imgs, int w, int ht, int nframes, int threads, int nfeatures,
                                   const float* world, const int* score, const double* xv13, const float* sigma2, int k,
                                   double sampleScale, int window, float nnratio, const double* cam9, int* out_counts,
                                   int* out_nmatch, int* out_sel, int* out_seln) {
    if (threads < 1) threads = 1;
    if (threads > nframes) threads = nframes;
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++)
        pool.emplace_back([=]() {
            const int s = (int)((long long)nframes * t / threads), e = (int)((long long)nframes * (t + 1) / threads);
            ORBextractor ex(nfeatures, 1.2f, 8, FAST_SCORE, 20, 7, 0, 0);
            Observability ob(cam9[0], (int)cam9[1], (int)cam9[2], cam9[3], cam9[4], cam9[5], cam9[6], cam9[7], cam9[8]);
            ob.rand_fn = &tl_rand;
            memcpy(ob.Xv, xv13, sizeof(ob.Xv));
            ob.predictPWLSVec(0.1, 1);
            std::vector<KeyPoint> kp[2]; std::vector<uint8_t> de[2];
            int cur = 0;
            bool havePrev = false;
            if (s > 0) { ex(imgs + (size_t)(s - 1) * w * ht, w, ht, w, kp[cur], de[cur]); havePrev = true; cur ^= 1; }
            std::vector<double> blocks, obsScore;
            std::vector<uint8_t> valid;
            std::vector<int> oct;
            for (int f = s; f < e; f++) {
                ex(imgs + (size_t)f * w * ht, w, ht, w, kp[cur], de[cur]);
                const int n = (int)kp[cur].size();
                out_counts[f] = n;
                int nm = 0;
                if (havePrev) {
                    Frame F1, F2;
                    F1.init(kp[cur ^ 1].data(), de[cur ^ 1].data(), (int)kp[cur ^ 1].size(), w, ht, 8, 1.2f);
                    F2.init(kp[cur].data(), de[cur].data(), n, w, ht, 8, 1.2f);
                    std::vector<float> pm(2 * kp[cur ^ 1].size());
                    for (size_t i = 0; i < kp[cur ^ 1].size(); i++) { pm[2 * i] = kp[cur ^ 1][i].x; pm[2 * i + 1] = kp[cur ^ 1][i].y; }
                    std::vector<int> m12;
                    ORBmatcher m(nnratio, true);
                    nm = m.SearchForInitialization(F1, F2, pm, m12, window);
                }
                out_nmatch[f] = nm;
                blocks.assign((size_t)n * 49, 0.0); obsScore.assign(n, 0.0); valid.assign(n, 1); oct.resize(n);
                for (int i = 0; i < n; i++) oct[i] = kp[cur][i].octave;
                ob.mKineIdx = 0;
                ob.batchInfoMat_Frame(world, oct.data(), score, valid.data(), n, sigma2, blocks.data(), obsScore.data(), nullptr);
                ob.lmkSelectPool.resize(n);
                for (int i = 0; i < n; i++) {
                    GoodPoint& g = ob.lmkSelectPool[i];
                    g.idx = i; g.obs_score = obsScore[i]; g.selected = false; g.upper_bound = 0;
                    memcpy(g.obs_block, &blocks[(size_t)49 * i], sizeof(g.obs_block));
                }
                tl_srand(1);
                std::vector<GoodPoint> sub;
                if (n > 0) ob.maxVolSelection_LazierGreedy(0, n, &sub, k, sampleScale, 1e30);
                out_seln[f] = (int)sub.size();
                for (size_t i = 0; i < sub.size() && (int)i < k; i++) out_sel[(size_t)f * k + i] = (int)sub[i].idx;
                havePrev = true;
                cur ^= 1;
            }
        });
    for (auto& th : pool) th.join();
    return 0;
}
