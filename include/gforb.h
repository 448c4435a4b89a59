/* gforb.h -- C ABI of the B200-native GF-ORB-SLAM tracking front end (libgforb.so).
 *
 * Drop-in boundary for ONE hot path of Aidenryan/GF_ORB_SLAM: ORB extraction -> Hamming matching -> good-feature
 * selection.  The reference has no FFI: the "operator API" is three C++ classes called synchronously from the
 * tracking thread.  Each entry point below names the reference member function it replaces (paths relative to the
 * reference tree); the headers under include/gforb/ re-create the class signatures on top of this ABI.
 *
 * Conventions
 *  - every function returns GFORB_OK (0) or a negative GFORB_ERR_*; nothing throws.
 *  - handles are opaque; one CUDA device per handle; work is enqueued on `stream` (a cudaStream_t passed as void*,
 *    NULL = the CUDA default stream) and, unless GFORB_ASYNC is set in `flags`, the call returns after the stream
 *    has been synchronised.
 *  - `flags & GFORB_IN_DEVICE` / `GFORB_OUT_DEVICE`: the input / output pointers are device pointers (no copies);
 *    otherwise they are host pointers and the call performs the H2D / D2H copies itself (pinned memory recommended).
 *  - batch-first: all frame-level calls take `batch` independent units.
 *  - There is no CPU fallback: if no CUDA device is usable every create() fails with GFORB_ERR_CUDA.
 */
#ifndef GFORB_H
#define GFORB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GFORB_OK 0
#define GFORB_ERR_CUDA (-1)        /* CUDA runtime error / no device */
#define GFORB_ERR_ARG (-2)         /* bad argument (the reference would assert) */
#define GFORB_ERR_CAPACITY (-3)    /* exceeds the sizes given at create() */
#define GFORB_ERR_GEOMETRY (-4)    /* cell grid degenerate for this image size (reference: cv::Exception / div by 0) */
#define GFORB_ERR_OVERFLOW (-5)    /* internal candidate list overflow (reported, never silent) */

#define GFORB_IN_DEVICE 1
#define GFORB_OUT_DEVICE 2
#define GFORB_ASYNC 4

/* Layout-identical to cv::KeyPoint (28 bytes): pt.x, pt.y, size, angle, response, octave, class_id. */
typedef struct GforbKeyPoint {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} GforbKeyPoint;

const char* gforb_version(void);
int gforb_device_count(void);

/* ------------------------------------------------------------------------------------------------ extraction
 * Replaces ORB_SLAM::ORBextractor (include/ORBextractor.h:51-104, src/ORBextractor.cc:464-998). */
typedef struct GforbExtractorConfig {
    int32_t nfeatures;        /* ctor arg nfeatures (1000) */
    float scaleFactor;        /* ctor arg scaleFactor (1.2f) */
    int32_t nlevels;          /* ctor arg nlevels (8) */
    int32_t scoreType;        /* 1 = FAST_SCORE, 0 = HARRIS_SCORE (src/ORBextractor.cc:86-127, :667-671) */
    int32_t iniThFAST;        /* ctor arg fastTh (20) */
    int32_t minThFAST;        /* hard-coded 7 in the reference (src/ORBextractor.cc:627) */
    int32_t blur_variant;     /* 0 = OpenCV>=3.4 bit-exact Gaussian (cv2-4.13 verified), 1 = legacy 2.4 integer kernel */
    int32_t retain_variant;   /* 0 = nth_element at begin()+n (OpenCV 2.4.13/3.4.1), 1 = begin()+n-1 */
    int32_t max_width, max_height, max_batch;
    int32_t device;           /* CUDA device ordinal */
} GforbExtractorConfig;

typedef struct GforbExtractor GforbExtractor;

int gforb_extractor_create(const GforbExtractorConfig* cfg, GforbExtractor** out);
int gforb_extractor_destroy(GforbExtractor* h);

/* ORBextractor::operator()(image, mask, keypoints, descriptors) for `batch` frames (src/ORBextractor.cc:769-875).
 * images: 8-bit single channel, row stride `stride`, frame b at images + b*frame_stride.  The mask argument of the
 * reference is dead code (always empty, src/Frame.cc:59) and is not part of the ABI.
 * keypoints: batch x nfeatures, descriptors: batch x nfeatures x 32 bytes, counts: batch (valid rows per frame);
 * output order is exactly the reference's (level-major, then cell order, then retainBest permutation). */
int gforb_extract(GforbExtractor* h, const uint8_t* images, int width, int height, int stride, size_t frame_stride,
                  int batch, GforbKeyPoint* keypoints, uint8_t* descriptors, int32_t* counts, int flags, void* stream);

/* ORBextractor::GetLevels / GetScaleFactor (include/ORBextractor.h:66-70) and the ctor tables
 * (mvScaleFactor, mnFeaturesPerLevel; src/ORBextractor.cc:469-494). */
int gforb_extractor_levels(const GforbExtractor* h);
float gforb_extractor_scale_factor(const GforbExtractor* h);
int gforb_extractor_tables(const GforbExtractor* h, float* scale_factors, int32_t* features_per_level);

/* Debug/parity taps on the last gforb_extract() call: copies level `level` of frame `b` (bordered plane,
 * (h+32) x pitch bytes, pitch returned) -- which = 0: unblurred pyramid (ComputePyramid output), 1: blurred ROI. */
int gforb_extractor_debug_level(GforbExtractor* h, int b, int level, int which, uint8_t* dst, size_t dst_bytes,
                                int* w, int* hgt, int* pitch);
/* number of kernels launched by the last gforb_extract() call on this handle */
int gforb_extractor_last_launches(const GforbExtractor* h);
/* name + device time (ms, CUDA events on the launch stream) of stage i of the last profiled call; returns the
 * number of stages.  Profiling is enabled with gforb_extractor_set_profiling(h, 1). */
int gforb_extractor_set_profiling(GforbExtractor* h, int on);
int gforb_extractor_stage_time(const GforbExtractor* h, int i, const char** name, float* ms);
/* 1 (default): within a call the Gaussian blur runs on the handle's own stream beside the latency-bound quota / retainBest /
 * orientation kernels and joins before the descriptors; 0: all kernels of a call in sequence on the caller's stream. */
int gforb_extractor_set_concurrency(GforbExtractor* h, int on);
/* Keypoint distribution: 0 (default) = the reference's cell quota (ComputeKeyPoints, src/ORBextractor.cc:535-758); 1 = the
 * quadtree of ORB-SLAM2 (ComputeKeyPointsOctTree / DistributeOctTree: FAST on ~30-px cells with "empty at iniThFAST ->
 * minThFAST", nodes split until mnFeaturesPerLevel, the best response of every node) -- the mode north_star names; the
 * GF-ORB-SLAM reference itself has no such code, so this mode is checked against the restatement in oracle/orb_quadtree.cpp
 * only (size-sort ties by node creation order, rows beyond nfeatures dropped).  FAST_SCORE only; re-plans the cell grid. */
int gforb_extractor_set_distribution(GforbExtractor* h, int mode);

/* ------------------------------------------------------------------------------------------------ matching
 * Replaces ORB_SLAM::ORBmatcher (include/ORBmatcher.h:39-294, src/ORBmatcher.cc) for the per-frame tracking
 * overloads, plus the Frame grid it searches (src/Frame.cc:118-131, 300-377).  MapPoint* / NULL is flattened to an
 * int32 id / -1.  Every call takes `batch` independent problems (frames / frame pairs).
 *
 * Pointer residency: with GFORB_IN_DEVICE all pointers inside the *Set structs and all pure-input arrays are device
 * pointers; with GFORB_OUT_DEVICE all output and in/out arrays are device pointers.  Otherwise they are host
 * pointers and the call stages them itself. */
typedef struct GforbFrameSet {
    const GforbKeyPoint* keypoints;   /* batch x kp_stride         Frame::mvKeysUn */
    const uint8_t* descriptors;       /* batch x kp_stride x 32    Frame::mDescriptors */
    const int32_t* counts;            /* batch                     Frame::N */
    int32_t kp_stride;
    int32_t min_x, max_x, min_y, max_y; /* Frame::mnMinX/mnMaxX/mnMinY/mnMaxY (src/Frame.cc:489-493) */
} GforbFrameSet;

typedef struct GforbMapPointSet {     /* the MapPoint fields SearchByProjection reads; arrays are batch x mp_stride */
    const float* proj_x;              /* MapPoint::mTrackProjX */
    const float* proj_y;              /* MapPoint::mTrackProjY */
    const int32_t* scale_level;       /* MapPoint::mnTrackScaleLevel */
    const float* view_cos;            /* MapPoint::mTrackViewCos */
    const uint8_t* flags;             /* bit0 = mbTrackInView, bit1 = isBad() */
    const uint8_t* descriptors;       /* MapPoint::GetDescriptor(), x 32 bytes */
    const int32_t* counts;            /* batch: vpMapPoints.size() */
    int32_t mp_stride;
} GforbMapPointSet;

typedef struct GforbMatcherConfig {
    int32_t max_batch;                /* problems per call */
    int32_t max_keypoints;            /* per frame (<= 16384) */
    int32_t max_queries;              /* per problem: map points (SearchByProjection) or source keypoints */
    int32_t cand_pool;                /* candidate entries per problem; 0 = 64 * max_queries */
    int32_t nlevels;                  /* Frame::mnScaleLevels */
    float scaleFactor;                /* Frame::mfScaleFactor -> mvScaleFactors[] */
    int32_t device;
} GforbMatcherConfig;

typedef struct GforbMatcher GforbMatcher;
int gforb_matcher_create(const GforbMatcherConfig* cfg, GforbMatcher** out);
int gforb_matcher_destroy(GforbMatcher* h);
/* kernels launched by the last call on this handle */
int gforb_matcher_last_launches(const GforbMatcher* h);

/* ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:2384-2400) for n descriptor pairs: out[i] = |a_i xor b_i|. */
int gforb_descriptor_distance(GforbMatcher* h, const uint8_t* a, const uint8_t* b, int n, int32_t* out, int flags,
                              void* stream);

/* Frame::GetFeaturesInArea (src/Frame.cc:300-365) on the grid the Frame ctor builds (:118-131, PosInGrid :367-377):
 * n_queries windows (x, y, r, minLevel, maxLevel) against frame `frame` of `frames`; out_idx receives, per query,
 * up to `cap` keypoint indices in the reference's order, out_n the full counts.  Parity tap for the grid. */
int gforb_features_in_area(GforbMatcher* h, const GforbFrameSet* frames, int batch, int frame, const float* x,
                           const float* y, const float* r, const int32_t* min_level, const int32_t* max_level,
                           int n_queries, int32_t* out_idx, int cap, int32_t* out_n, int flags, void* stream);

/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (src/ORBmatcher.cc:1172-1287).
 * prev_matched: in/out, batch x F1.kp_stride x 2;  matches12: out, batch x F1.kp_stride;  nmatches: out, batch. */
int gforb_search_for_initialization(GforbMatcher* h, const GforbFrameSet* F1, const GforbFrameSet* F2, int batch,
                                    float* prev_matched, int32_t* matches12, int32_t* nmatches, int window_size,
                                    float nnratio, int check_orientation, int flags, void* stream);

/* ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>&, th) (src/ORBmatcher.cc:384-465).
 * frame_mappoints / match_score: in/out, batch x F.kp_stride (Frame::mvpMapPoints as ids, Frame::mvpMatchScore);
 * nmatches: out, batch. */
int gforb_search_by_projection(GforbMatcher* h, const GforbFrameSet* F, const GforbMapPointSet* mps, int batch,
                               float th, float nnratio, int32_t* frame_mappoints, int32_t* match_score,
                               int32_t* nmatches, int flags, void* stream);

/* ORBmatcher::SearchByProjection_GoodFeature(Frame& F, const vector<MapPoint*>&, th_good, th_rest)
 * (src/ORBmatcher.cc:51-140): as gforb_search_by_projection, but the search window of a map point with
 * goodAtFrameId == F.mnId (bit2 of GforbMapPointSet::flags) is scaled by th_good and of the others by th_rest (when
 * th_rest != 1); Frame::mvpMatchScore is not written by this overload.
 * ORBmatcher::SearchByProjection_Budget(F, MPs, th, time_constr) (:276-379) is gforb_search_by_projection when
 * time_constr > 0 and matches nothing otherwise (wall-clock budgets are unlimited on this path). */
int gforb_search_by_projection_good_feature(GforbMatcher* h, const GforbFrameSet* F, const GforbMapPointSet* mps, int batch,
                                            float th_good, float th_rest, float nnratio, int32_t* frame_mappoints,
                                            int32_t* nmatches, int flags, void* stream);

/* MapPoint::ComputeDistinctiveDescriptors() (src/MapPoint.cc:196-262) for n_points map points: descriptors = the
 * observation descriptors (KeyFrame::GetDescriptor of every non-bad observing keyframe) of all points concatenated,
 * offsets[i] .. offsets[i+1] = the rows of point i (n_points + 1 entries).  best_idx[i] = the row (relative to
 * offsets[i]) with the least median Hamming distance to the point's other observations, -1 for a point without
 * observations (the reference returns and keeps mDescriptor); out_desc (n_points x 32, may be NULL) receives that
 * descriptor (= the new MapPoint::mDescriptor; rows of empty points are left as they are). */
int gforb_mappoint_distinctive_descriptors(GforbMatcher* h, int n_points, const uint8_t* descriptors, const int32_t* offsets,
                                           int32_t* best_idx, uint8_t* out_desc, int flags, void* stream);

/* Frame::isInFrustum(MapPoint*, viewingCosLimit) (src/Frame.cc:166-227) for every map point of `batch` frames: the
 * pass that produces the GforbMapPointSet fields SearchByProjection reads.  Rcw / tcw / Ow: batch x 9 / 3 / 3 (Frame::mRcw
 * row-major, mtcw, mOw); world / normal: batch x mp_stride x 3 (MapPoint::GetWorldPos / GetNormal); min_distance /
 * max_distance: MapPoint::GetMin/MaxDistanceInvariance.  proj_x, proj_y, scale_level, view_cos are written for the
 * points in view; mp_flags: in/out, bit0 <- mbTrackInView, the other bits are kept. */
int gforb_frame_is_in_frustum(GforbMatcher* h, int batch, const float* Rcw, const float* tcw, const float* Ow, float fx,
                              float fy, float cx, float cy, int min_x, int max_x, int min_y, int max_y, const float* world,
                              const float* normal, const float* min_distance, const float* max_distance,
                              const int32_t* counts, int mp_stride, float viewing_cos_limit, float* proj_x, float* proj_y,
                              int32_t* scale_level, float* view_cos, uint8_t* mp_flags, int flags, void* stream);

/* Frame::UndistortKeyPoints() (src/Frame.cc:389-423) for the keypoints of `batch` frames: keys_un[i] = keys[i] with pt
 * replaced by cv::undistortPoints(pt, K, dist_coef, R = I, P = K); dist_coef: HOST array of n_dist (0, 4, 5 or 8) floats
 * = Frame::mDistCoef (k1, k2, p1, p2[, k3[, k4, k5, k6]]); dist_coef[0] == 0 copies the keypoints (the reference's early
 * return).  Only the first counts[b] entries of every frame are written. */
int gforb_frame_undistort_keypoints(GforbMatcher* h, int batch, const GforbKeyPoint* keys, const int32_t* counts,
                                    int kp_stride, float fx, float fy, float cx, float cy, const float* dist_coef,
                                    int n_dist, GforbKeyPoint* keys_un, int flags, void* stream);

/* ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th) (src/ORBmatcher.cc:2081-2202).
 * last_mappoints: batch x last.kp_stride ids (-1 = NULL); last_outlier: same shape or NULL; last_world: batch x
 * last.kp_stride x 3 (MapPoint::GetWorldPos of LastFrame.mvpMapPoints[i]); Tcw: batch x 16 row-major (Current.mTcw);
 * cur_mappoints / cur_score: in/out, batch x cur.kp_stride. */
int gforb_search_by_projection_frames(GforbMatcher* h, const GforbFrameSet* cur, const GforbFrameSet* last, int batch,
                                      const int32_t* last_mappoints, const uint8_t* last_outlier,
                                      const float* last_world, const float* Tcw, float fx, float fy, float cx, float cy,
                                      float th, float nnratio, int check_orientation, int32_t* cur_mappoints,
                                      int32_t* cur_score, int32_t* nmatches, int flags, void* stream);

/* ORBmatcher::SearchByProjection_GoodFeature(Frame& Current, const Frame& Last, th_good, th_rest)
 * (src/ORBmatcher.cc:143-274; the call Tracking::TrackWithMotionModel makes, src/Tracking.cc:1442): as
 * gforb_search_by_projection_frames with the radius th_good * scale for last-frame map points with
 * goodAtFrameId == Current.mnId (last_good: batch x last.kp_stride, non-zero = good) and th_rest * scale for the others;
 * no match score is stored or reset. */
int gforb_search_by_projection_frames_good_feature(GforbMatcher* h, const GforbFrameSet* cur, const GforbFrameSet* last, int batch,
                                                   const int32_t* last_mappoints, const uint8_t* last_outlier,
                                                   const uint8_t* last_good, const float* last_world, const float* Tcw, float fx,
                                                   float fy, float cx, float cy, float th_good, float th_rest, float nnratio,
                                                   int check_orientation, int32_t* cur_mappoints, int32_t* nmatches, int flags,
                                                   void* stream);

/* ORBmatcher::WindowSearch(F1, F2, windowSize, vpMapPointMatches2, minScaleLevel, maxScaleLevel)
 * (src/ORBmatcher.cc:979-1086).  f1_mappoints: batch x F1.kp_stride ids (>= 0: non-NULL, not bad);
 * matches2: out, batch x F2.kp_stride (map-point id or -1). */
int gforb_window_search(GforbMatcher* h, const GforbFrameSet* F1, const GforbFrameSet* F2, int batch,
                        const int32_t* f1_mappoints, int window_size, int min_scale_level, int max_scale_level,
                        float nnratio, int check_orientation, int32_t* matches2, int32_t* nmatches, int flags,
                        void* stream);

/* ORBmatcher::SearchByProjection(Frame& F1, Frame& F2, int windowSize, vector<MapPoint*>& vpMapPointMatches2)
 * (src/ORBmatcher.cc:1089-1168; Tracking::TrackPreviousFrame, src/Tracking.cc:1373-1376): F1 map points that F2 does not
 * hold yet, projected with F2's pose into a fixed pixel window on the same pyramid level; ratio test
 * best <= second * nnratio and best <= TH_HIGH.  f1_mappoints: batch x F1.kp_stride ids (>= 0: non-NULL, not bad);
 * f1_world: batch x F1.kp_stride x 3 (MapPoint::GetWorldPos); Tcw2: batch x 16 row-major (F2.mTcw);
 * matches2: in/out, batch x F2.kp_stride, in = F2.mvpMapPoints, out = vpMapPointMatches2; f2_score: in/out F2.mvpMatchScore. */
int gforb_search_by_projection_window(GforbMatcher* h, const GforbFrameSet* F1, const GforbFrameSet* F2, int batch,
                                      const int32_t* f1_mappoints, const float* f1_world, const float* Tcw2, float fx, float fy,
                                      float cx, float cy, int window_size, float nnratio, int32_t* matches2, int32_t* f2_score,
                                      int32_t* nmatches, int flags, void* stream);

/* ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches) (src/ORBmatcher.cc:724-853;
 * Tracking::TrackReferenceKeyFrame and Relocalization).  The vocabulary transform is the caller's (DBoW2): kf_nodes /
 * f_nodes = the FeatureVector node id of every keyframe / frame feature (batch x kp_stride, -1 = in no node).  The walk is
 * the reference's: nodes ascending, keyframe features of a node in index order, every frame feature of the same node that
 * is still unmatched as candidate, best <= TH_LOW and best < nnratio * second, rotation histogram under check_orientation.
 * kf_mappoints: batch x KF.kp_stride ids of pKF->GetMapPointMatches() (-1 = NULL or isBad()); KF->keypoints = the
 * keyframe's undistorted keys (angle), KF->descriptors its descriptors.  matches: out, batch x F.kp_stride (id or -1). */
int gforb_search_by_bow(GforbMatcher* h, const GforbFrameSet* KF, const int32_t* kf_mappoints, const int32_t* kf_nodes,
                        const GforbFrameSet* F, const int32_t* f_nodes, int batch, float nnratio, int check_orientation,
                        int32_t* matches, int32_t* nmatches, int flags, void* stream);

/* ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound, float th,
 * int ORBdist) (src/ORBmatcher.cc:2204-2336; Tracking::Relocalization).  Per keyframe keypoint (batch x KF.kp_stride):
 * kf_mappoints = id of pKF->GetMapPointMatches()[i] (-1 = NULL), kf_skip (may be NULL) non-zero = isBad() or in
 * sAlreadyFound, kf_world (x 3) / kf_min_distance / kf_mp_descriptors (x 32) = GetWorldPos / GetMinDistanceInvariance /
 * GetDescriptor of that map point; KF->keypoints = the keyframe's undistorted keys (angle).  Tcw: batch x 16 (Current.mTcw);
 * the level is predicted from |x3Dw - Ow| / minDistance, the window is th * scale[level] over levels level-1 .. level+1,
 * the best unclaimed candidate is taken if its distance is <= orb_dist.  cur_mappoints / cur_score: in/out. */
int gforb_search_by_projection_keyframe(GforbMatcher* h, const GforbFrameSet* cur, const GforbFrameSet* KF, int batch,
                                        const int32_t* kf_mappoints, const uint8_t* kf_skip, const float* kf_world,
                                        const float* kf_min_distance, const uint8_t* kf_mp_descriptors, const float* Tcw,
                                        float fx, float fy, float cx, float cy, float th, int orb_dist, int check_orientation,
                                        int32_t* cur_mappoints, int32_t* cur_score, int32_t* nmatches, int flags, void* stream);

/* ---- the ORBmatcher overloads of the LocalMapping / LoopClosing threads: same Hamming core, candidate
 * order and claim rules as above.  Map mutations (MapPoint::Replace, AddObservation, KeyFrame::AddMapPoint) stay with the
 * caller: the calls return what the reference's matching loop decides, the caller applies it to its map. */

/* ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12) (src/ORBmatcher.cc:1289-1424;
 * LoopClosing::ComputeSim3).  kf*_mappoints: batch x kp_stride ids of GetMapPointMatches() (-1 = NULL or isBad());
 * kf*_nodes: DBoW2 FeatureVector node of every feature (-1 = none).  Walk as in gforb_search_by_bow; candidates are the
 * KF2 features of the node that hold a map point and are not matched yet; best < TH_LOW (strict) and best < nnratio *
 * second.  matches12: out, batch x KF1.kp_stride = the id vpMapPoints2[bestIdx2] (or -1). */
int gforb_search_by_bow_keyframes(GforbMatcher* h, const GforbFrameSet* KF1, const int32_t* kf1_mappoints, const int32_t* kf1_nodes,
                                  const GforbFrameSet* KF2, const int32_t* kf2_mappoints, const int32_t* kf2_nodes, int batch,
                                  float nnratio, int check_orientation, int32_t* matches12, int32_t* nmatches, int flags,
                                  void* stream);

/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedKeys1, vMatchedKeys2, vMatchedPairs) (src/ORBmatcher.cc:1426-1588;
 * LocalMapping::CreateNewMapPoints) with CheckDistEpipolarLine (:705-722).  Queries: KF1 features WITHOUT a map point
 * (kf1_mappoints < 0); candidates: KF2 features of the same node without a map point, not matched yet, distance <= TH_LOW,
 * visited in (distance, index) order while distance <= 2 * best; the first one within 3.84 sigma2(octave) of the epipolar
 * line of F12 (batch x 9, row-major) is taken.  matches12: out, batch x KF1.kp_stride = index into KF2 or -1 (the
 * reference's vMatchedPairs are the pairs (i, matches12[i]) in ascending i). */
int gforb_search_for_triangulation(GforbMatcher* h, const GforbFrameSet* KF1, const int32_t* kf1_mappoints, const int32_t* kf1_nodes,
                                   const GforbFrameSet* KF2, const int32_t* kf2_mappoints, const int32_t* kf2_nodes, int batch,
                                   const float* F12, int check_orientation, int32_t* matches12, int32_t* nmatches, int flags,
                                   void* stream);

typedef struct GforbWorldPointSet {   /* MapPoint fields the projection overloads below read; arrays are batch x mp_stride */
    const float* world;               /* x 3: GetWorldPos() */
    const float* normal;              /* x 3: GetNormal() (not read by gforb_search_by_sim3; may be NULL there) */
    const float* min_distance;        /* GetMinDistanceInvariance() */
    const float* max_distance;        /* GetMaxDistanceInvariance() */
    const uint8_t* descriptors;       /* x 32: GetDescriptor() */
    const uint8_t* skip;              /* non-zero: the reference `continue`s before projecting (NULL, isBad(), IsInKeyFrame /
                                         already found / already matched) */
    const int32_t* counts;            /* batch: list length */
    int32_t mp_stride;
} GforbWorldPointSet;

/* The matching loop of ORBmatcher::Fuse(KeyFrame* pKF, vector<MapPoint*>& vpMapPoints, float th) (src/ORBmatcher.cc:1590-1708,
 * is_sim3 = 0, pose = Tcw, batch x 16 row-major; Ow = -Rcw.t()*tcw as KeyFrame::SetPose computes it) and of
 * Fuse(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, float th) (:1710-1839, is_sim3 = 1, pose = Scw,
 * decomposed as :1718-1723).  Per list entry: projection, KeyFrame::IsInImage, distance / viewing-angle gates, predicted level,
 * GetFeaturesInArea(u, v, th * scale[level]), levels level-1 .. level, first smallest distance (no claims).  best_idx /
 * best_dist: out, batch x mp_stride (-1 / INT_MAX where the point was skipped or had no candidate); nfused: out, batch =
 * entries with best_dist <= TH_LOW.  The caller walks the list in order and applies Replace / AddObservation (a point that
 * an earlier entry made bad or added to the keyframe is skipped there, as in the reference). */
int gforb_fuse(GforbMatcher* h, const GforbFrameSet* KF, int batch, const float* pose, int is_sim3, float fx, float fy, float cx,
               float cy, const GforbWorldPointSet* points, float th, int32_t* best_idx, int32_t* best_dist, int32_t* nfused,
               int flags, void* stream);

/* ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, vector<MapPoint*>& vpMatched,
 * int th) (src/ORBmatcher.cc:855-977; LoopClosing::ComputeSim3 / CorrectLoop).  points->skip = isBad() or in vpMatched at
 * entry; point_ids: batch x mp_stride, the id written for a match; matched: in/out, batch x KF.kp_stride (vpMatched as ids,
 * -1 = NULL): keypoints that hold an id are skipped, best <= TH_LOW claims the keypoint. */
int gforb_search_by_projection_sim3(GforbMatcher* h, const GforbFrameSet* KF, int batch, const float* Scw, float fx, float fy,
                                    float cx, float cy, const GforbWorldPointSet* points, const int32_t* point_ids, float th,
                                    int32_t* matched, int32_t* nmatches, int flags, void* stream);

/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) (src/ORBmatcher.cc:1841-2079; LoopClosing::ComputeSim3).
 * points1 / points2: the map points of KF1 / KF2 per keypoint (mp_stride = kp_stride; skip = NULL, isBad() or
 * vbAlreadyMatched1/2); T1w / T2w: batch x 16 (pKF->GetRotation / GetTranslation); s12: batch; R12: batch x 9; t12: batch x 3;
 * fx.. = pKF1's intrinsics (the reference uses them for both directions).  Both directions are searched without claims
 * (best <= TH_HIGH), a pair is kept when the two agree.  kf2_mappoints: ids of KF2's map points; matches12: in/out, batch x
 * KF1.kp_stride (agreeing pairs receive kf2_mappoints[idx2]); nfound: out, batch. */
int gforb_search_by_sim3(GforbMatcher* h, const GforbFrameSet* KF1, const GforbFrameSet* KF2, int batch,
                         const GforbWorldPointSet* points1, const GforbWorldPointSet* points2, const int32_t* kf2_mappoints,
                         const float* T1w, const float* T2w, const float* s12, const float* R12, const float* t12, float fx, float fy,
                         float cx, float cy, float th, int32_t* matches12, int32_t* nfound, int flags, void* stream);

/* Batched all-pairs Hamming top-2 of n_q query descriptors against one database shard of n_db descriptors
 * (BASELINE.json config 5; the reference's brute-force analogue is the DescriptorDistance loop of
 * MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:225-250).  out: n_q x 2 packed keys
 * (dist << 32 | global index), best then second; ties -> lower index; idx_base = global index of db row 0.
 * The cross-shard merge is gforb_top2_merge over the all-gathered keys (NCCL all-gather by the host layer). */
int gforb_allpairs_top2(GforbMatcher* h, const uint8_t* queries, int n_q, const uint8_t* db, int64_t n_db,
                        int64_t idx_base, uint64_t* out_keys, int flags, void* stream);
int gforb_top2_merge(GforbMatcher* h, const uint64_t* gathered /* n_shards x n_q x 2 */, int n_shards, int n_q,
                     uint64_t* out_keys /* n_q x 2 */, int flags, void* stream);

/* Allocation-free variant of gforb_allpairs_top2 (tensor-core kernel).  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  No handle and no allocation of any kind: the per-split partial keys and the kernel's
 * error word live in a file-scope device array of 1 MiB + 256 B that every GPU loading the library holds until exit.
 * Limits: n_q <= 4096 and (database splits) x n_q <= 65536 key pairs, where splits = ceil(tiles / min(ceil(tiles / max(1,
 * SMs / ceil(n_q / 128))), 8192)) and tiles = ceil(n_db / 256) -- on a 148-SM B200 every n_db <= 2^24 fits at any n_q <= 4096;
 * beyond that the call returns GFORB_ERR_CAPACITY (and says why on stderr).  queries / db / out_keys: device pointers
 * (flags must hold GFORB_IN_DEVICE | GFORB_OUT_DEVICE; GFORB_ASYNC skips the synchronise + error read), same layouts and the
 * same keys as gforb_allpairs_top2.  Every call on a device shares the one array, which keeps the last call's contents (the
 * error word is cleared on the stream at the start of each call): calls that overlap on the GPU race -- issue them on one
 * stream or order them with events.  The first call on a device does the one-time set-up (cudaGetSymbolAddress, kernel
 * attribute); make it before starting a capture. */
int gforb_allpairs_top2_static(int device, const uint8_t* queries, int n_q, const uint8_t* db, int64_t n_db, int64_t idx_base,
                               uint64_t* out_keys, int flags, void* stream);
/* gforb_allpairs_top2 runs on the tensor cores (tcgen05.mma kind::i8 on +-1 / 0-1 expanded descriptors, exact); the
 * environment variable GFORB_ALLPAIRS=popc selects the integer-pipe (LOP3/POPC) kernel, the parity reference. */

/* Calls made with GFORB_ASYNC do not read the handle's device error word (candidate-pool overflow etc. -- the reference
 * has no such limit, so an overflow would silently change results).  gforb_matcher_poll_error synchronises `stream` and
 * returns GFORB_OK or the first GFORB_ERR_* recorded since the last poll (and clears it). */
int gforb_matcher_poll_error(GforbMatcher* h, void* stream);

/* ------------------------------------------------------------------------------------------------ good features
 * Replaces ORB_SLAM::Observability (include/Observability.h:113-770, include/Util.hpp, src/Observability.cc) for
 * DIMENSION_OF_STATE_MODEL = 7: kinematics (host), per-match information blocks and log-det subset selection. */
typedef struct GforbObsConfig {
    double fu, fv;                 /* PinHoleCamera (include/Util.hpp:131-167): ctor (f,...,dx,dy) passes fu = f/dx */
    int32_t nRows, nCols;
    double Cx, Cy, k1, k2, dx, dy;
    int32_t max_batch;             /* frames / selection problems per call */
    int32_t max_points;            /* matched map points per frame (<= 8192) */
    int32_t device;
} GforbObsConfig;

typedef struct GforbObs GforbObs;
int gforb_obs_create(const GforbObsConfig* cfg, GforbObs** out);
int gforb_obs_destroy(GforbObs* h);
int gforb_obs_last_launches(const GforbObs* h);
/* Scheduling hint for gforb_obs_select's lazier-greedy kernel: register budget for 2 resident CTAs per SM (168 registers, no
 * spills: fastest when the selection has the GPU to itself), for 3 (128 registers: a third less of the register file, better
 * when it runs beside other kernels on another stream), or 0 = chosen from the batch size.  Results do not depend on it. */
int gforb_obs_set_select_residency(GforbObs* h, int ctas_per_sm);

/* Observability::Xv (13-state [r(3) q(4) v(3) w(3)]), updatePWLSVec (include/Observability.h:222-259) and
 * predictPWLSVec (:261-295); matrices are 4x4 row-major float like the reference's cv::Mat.  Host-side. */
int gforb_obs_set_state(GforbObs* h, const double* Xv13);
int gforb_obs_get_state(const GforbObs* h, double* Xv13);
int gforb_obs_update_pwls(GforbObs* h, double t_prev, const float* Tcw_prev, double t_cur, const float* Twc_cur);
int gforb_obs_predict_pwls(GforbObs* h, double dt, int num_seg);
int gforb_obs_num_segments(const GforbObs* h);
/* kinematic[seg] (KineStruct, include/Observability.h:94-110); any output pointer may be NULL */
int gforb_obs_get_segment(const GforbObs* h, int seg, double* Xv13, double* F169, double* F_Q16, double* F_Omg12,
                          double* F_Q_inSeg16, double* F_Omg_inSeg12, float* Tcw16);

/* batchInfoMat_Frame (src/Observability.cc:386-554) for `batch` frames: per matched point the 2x7 Jacobian
 * (compute_H_subblock_simplied), octave re-weighting (reWeightInfoMat) and the 7x7 block H^T H.
 * Xv: batch x 13 (kinematic[mKineIdx].Xv of each frame); world: batch x stride x 3 (MapPoint::GetWorldPos);
 * valid = pMP && !mvbOutlier && mvbCandidate; blocks: batch x stride x 49 row-major; obs_score: ObsScore (match
 * score, -1 when not valid); H_meas: batch x stride x 14 or NULL. */
int gforb_obs_build_info_frame(GforbObs* h, const double* Xv, const float* world, const int32_t* octave,
                               const int32_t* match_score, const uint8_t* valid, const int32_t* counts, int stride,
                               int batch, const float* level_sigma2, int nlevels, double* blocks, double* obs_score,
                               double* H_meas, int flags, void* stream);
/* batchInfoMat_Map (src/Observability.cc:556-644): identity noise, optional visibility check; obs_score is 1 (built),
 * -1 (not visible) or -2 (skipped); uv: batch x stride x 2 projections (MapPoint::u_proj, v_proj) or NULL;
 * H_meas: batch x stride x 14 (MapPoint::H_meas = [H13 H47], read back by gforb_obs_active_match) or NULL. */
int gforb_obs_build_info_map(GforbObs* h, const double* Xv, const float* world, const uint8_t* track_in_view,
                             const int32_t* counts, int stride, int batch, int check_viz, int min_x, int max_x,
                             int min_y, int max_y, double* blocks, double* obs_score, float* uv, double* H_meas,
                             int flags, void* stream);
/* logDet (include/Util.hpp:714-731) of n 7x7 row-major matrices. */
int gforb_obs_logdet(GforbObs* h, const double* mats, int n, double* out, int flags, void* stream);
/* the device's glibc-TYPE_3 rand() stream (parity tap; out_host is a host pointer) */
int gforb_obs_rand_stream(GforbObs* h, uint32_t seed, int n, int32_t* out_host);

/* Subset selection over the pool lmkSelectPool = entries 0..counts[b]-1 of problem b (src/Observability.cc):
 * mtd 1 = maxVolSelection_BaselineGreedy (:3031), 2 = maxVolSelection_LazierGreedy (:2815),
 * 3 = maxVolAutomatic_LazierGreedy (:3141), 5 = maxVolDeletion_LazierGreedy (:2339); k = mpLimit.
 * The reference draws from the process-global rand(); here the stream is an input: seeds[b] (NULL = 1, the state of a
 * process that never called srand) restarts glibc's generator for problem b.  Time caps are infinite.
 * out_idx: batch x out_stride pool indices (selection order; deletion: ascending), out_logdet: running log-det after
 * each pick (selection modes) or NULL, out_n: picks per problem. */
int gforb_obs_select(GforbObs* h, const double* blocks, const double* obs_score, const int32_t* counts, int stride,
                     int batch, int k, int mtd, double sample_scale, const uint32_t* seeds, int32_t* out_idx,
                     double* out_logdet, int32_t* out_n, int out_stride, int flags, void* stream);

/* Observability::runActiveMapMatching(Frame*, mat_type, mBaseInfoMat, th, ORBmatcher&, num_to_match, time)
 * (src/Observability.cc:1249-1524) for `batch` frames: greedy log-det ranking of the visible local-map points
 * interleaved with ORBmatcher::SearchByProjection_OnePoint (include/ORBmatcher.h:71-145).
 * mps: Observability::mMapPoints flattened; obs_score (batch x mp_stride, from gforb_obs_build_info_map; > 0 means
 * MapPoint::updateAtFrameId == pFrame->mnId; NULL = every point), obs_mat = MapPoint::ObsMat (x 49), H_meas =
 * MapPoint::H_meas (x 14), base_info = mBaseInfoMat (batch x 49), level_sigma2 = Frame::mvLevelSigma2.
 * mat_type: 0 = MAP_INFO_MATRIX, 1 = FRAME_INFO_MATRIX (the hybrid types need DIMENSION_OF_STATE_MODEL = 13).
 * frame_mappoints / match_score: in/out as in gforb_search_by_projection; nmatched: return value per frame;
 * left_idx (batch x mp_stride) / left_n: mLeftMapPoints as indices into mps; out_info: batch x 49 accumulated
 * matrix or NULL.  rand(): seeds[b] restarts glibc's generator per frame (NULL = 1); the time cap is infinite. */
int gforb_obs_active_match(GforbObs* obs, GforbMatcher* matcher, const GforbFrameSet* F, const GforbMapPointSet* mps,
                           int batch, const double* obs_score, const double* obs_mat, const double* H_meas,
                           const double* base_info, const float* level_sigma2, int nlevels, int mat_type, float th,
                           float nnratio, int num_to_match, const uint32_t* seeds, int32_t* frame_mappoints,
                           int32_t* match_score, int32_t* nmatched, int32_t* left_idx, int32_t* left_n,
                           double* out_info, int flags, void* stream);

/* ------------------------------------------------------------------------------------------------ pose optimisation
 * Optimizer::PoseOptimization(Frame*) (src/Optimizer.cc:279-413; the step right after the matchers of this path:
 * TrackWithMotionModel :1596, TrackLocalMap :2776) for `batch` frames: g2o Levenberg-Marquardt on one SE3 vertex per
 * frame, EdgeSE3ProjectXYZ edges with Huber delta sqrt(5.991), information mvInvLevelSigma2[octave], 4 rounds of
 * {10,10,7,5} iterations with the chi2 outlier re-test {9.210, 7.378, 5.991, 5.991}.
 * Tcw: in/out, batch x 16 row-major float (Frame::mTcw); keys_un: batch x stride (Frame::mvKeysUn: pt and octave are read);
 * map_points: batch x stride ids (>= 0: Frame::mvpMapPoints[i] != NULL); world: batch x stride x 3 (MapPoint::GetWorldPos);
 * inv_level_sigma2: HOST array of nlevels floats (Frame::mvInvLevelSigma2); outlier: in/out batch x stride
 * (Frame::mvbOutlier; only entries with a map point are written); n_inliers: batch (the return value).
 * GFORB_IN_DEVICE covers keys_un / map_points / world / counts, GFORB_OUT_DEVICE covers Tcw / outlier / n_inliers. */
typedef struct GforbPoseOpt GforbPoseOpt;
int gforb_poseopt_create(int max_batch, int max_points, int device, GforbPoseOpt** out);
int gforb_poseopt_destroy(GforbPoseOpt* h);
int gforb_pose_optimization(GforbPoseOpt* h, float* Tcw, const GforbKeyPoint* keys_un, const int32_t* map_points, const float* world,
                            const int32_t* counts, int stride, int batch, const float* inv_level_sigma2, int nlevels, float fx,
                            float fy, float cx, float cy, uint8_t* outlier, int32_t* n_inliers, int flags, void* stream);

/* ------------------------------------------------------------------------------------------------ pipelined front end
 * The per-frame chain of Tracking::GrabImage (src/Tracking.cc:461-917) for a batch of consecutive frames, scheduled
 * below the ABI in C++ (CUDA streams / events; double-buffered so that step t+1's upload and extraction overlap step t's
 * matching, selection and result download):
 *   Frame ctor:  ORBextractor::operator()                                  (src/Frame.cc:59)
 *   tracking:    ORBmatcher::SearchForInitialization(frame t-1, frame t)    (src/Tracking.cc:1008,1097; the map-free matcher)
 *   good feats:  Observability::runMatrixBuilding(FRAME_INFO_MATRIX) + maxVolSelection_LazierGreedy, which -- as in the
 *                reference, where they work on the frame's matched map points -- start only after the matcher has finished.
 * Frame t-1 of a step's first frame is the last frame of the previous step (1-frame halo; the very first frame of a
 * stream is matched against an empty frame).  Every call is asynchronous; gforb_frontend_sync() joins all streams and
 * returns the first error any kernel recorded (the device error words of the three operators are polled there). */
typedef struct GforbFrontEndConfig {
    int32_t width, height;            /* frame size (8-bit, tightly packed rows) */
    int32_t nfeatures, nlevels;       /* ORBextractor ctor */
    float scaleFactor;
    int32_t iniThFAST, minThFAST;
    int32_t max_batch;                /* frames per step */
    int32_t window_size;              /* SearchForInitialization windowSize (100) */
    float nnratio;                    /* ORBmatcher ctor (0.9) */
    int32_t check_orientation;
    int32_t n_good;                   /* k of the good-feature subset (100) */
    double sample_scale;              /* lazier-greedy sampleScale (setSelction_Number :826) */
    double fu, fv, Cx, Cy, k1, k2, dx, dy;   /* Observability / PinHoleCamera */
    int32_t overlap;                  /* 1: matcher and selection on their own streams; 0: everything on one stream */
    int32_t device;
} GforbFrontEndConfig;

typedef struct GforbFrontEndOutputs {  /* per step; host pointers (pinned recommended) for step_host, device pointers from results() */
    GforbKeyPoint* keypoints;         /* batch x nfeatures */
    uint8_t* descriptors;             /* batch x nfeatures x 32 */
    int32_t* counts;                  /* batch */
    int32_t* matches12;               /* batch x nfeatures: vnMatches12 of (frame t-1 -> frame t) */
    int32_t* nmatches;                /* batch */
    int32_t* good_idx;                /* batch x n_good: selected keypoint indices, selection order */
    double* good_logdet;              /* batch x n_good or NULL */
    int32_t* good_n;                  /* batch */
} GforbFrontEndOutputs;

typedef struct GforbFrontEnd GforbFrontEnd;
int gforb_frontend_create(const GforbFrontEndConfig* cfg, GforbFrontEnd** out);
int gforb_frontend_destroy(GforbFrontEnd* h);
/* The matched map points the information blocks are built from, one per keypoint slot (MapPoint::GetWorldPos,
 * Frame::mvpMatchScore), the 13-state Xv of the predicted segment and Frame::mvLevelSigma2; host arrays. */
int gforb_frontend_set_map(GforbFrontEnd* h, const float* world /* nfeatures x 3 */, const int32_t* match_score /* nfeatures */,
                           const double* Xv13, const float* level_sigma2 /* nlevels */);
/* One step over `batch` frames resident in HBM (device pointer, batch x height x width). */
int gforb_frontend_step_device(GforbFrontEnd* h, const uint8_t* d_images, int batch);
/* One step over `batch` frames in host memory: H2D of the frames, the step, D2H of every output into *out (all inside
 * the call's stream work; out's arrays must stay valid until gforb_frontend_sync). */
int gforb_frontend_step_host(GforbFrontEnd* h, const uint8_t* images, int batch, const GforbFrontEndOutputs* out);
int gforb_frontend_sync(GforbFrontEnd* h);
/* Host-side wait for one gforb_frontend_step_host call only: the call issued `steps_back` (0 .. 6) calls before the most
 * recent one has read its host frames and delivered its host outputs (a caller cycling through S sets of staging buffers
 * waits with steps_back = S - 1 before refilling a set).  Does not poll the error words. */
int gforb_frontend_wait_results(GforbFrontEnd* h, int steps_back);
/* device pointers of the most recently issued step's results (valid after gforb_frontend_sync) */
int gforb_frontend_results(GforbFrontEnd* h, GforbFrontEndOutputs* device_out);
int gforb_frontend_launches_per_step(const GforbFrontEnd* h);
/* profiling: CUDA-event time of stage i ("extract", "match", "info+select", then the extractor's own stages) averaged over
 * the steps since profiling was switched on; returns the number of stages */
int gforb_frontend_set_profiling(GforbFrontEnd* h, int on);
int gforb_frontend_stage_time(GforbFrontEnd* h, int i, const char** name, float* ms);
/* error words of asynchronous calls on the other handles (see gforb_matcher_poll_error) */
int gforb_extractor_poll_error(GforbExtractor* h, void* stream);
int gforb_obs_poll_error(GforbObs* h, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GFORB_H */
