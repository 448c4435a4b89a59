// TEST INFRASTRUCTURE ONLY -- CPU oracle for the GF-ORB-SLAM tracking front end.
//
// This directory is a plain C++17 restatement (no OpenCV / Armadillo / ROS) of the reference's
// CPU algorithm for the hot path named in BASELINE.json.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may load it; the product library
// (gf_orb_slam_b200/csrc -> libgforb.so) never links, includes or calls anything in here.
//
// Parity pinning status (see DESIGN.md section "Oracle"):
//   * Observability kinematics / Jacobians: pinned against the reference's own golden vectors
//     (test/test_Kine_1.cpp:121-183, test/test_Jacobian.cpp:119-289) in tests/test_oracle_obs.py.
//   * OpenCV primitives used by ORBextractor (resize, copyMakeBorder, FAST, GaussianBlur,
//     fastAtan2): the reference's tests hold no vectors; pinned against cv2 4.13 (the only
//     OpenCV importable here; reference pins 2.4.13/3.4.1) in tests/test_oracle_cv.py and by
//     committed fixtures in tests/golden/.  => extraction/matching parity is "unpinned" w.r.t.
//     the reference's own tests, pinned w.r.t. OpenCV 4.13 primitives.
//
// All file:line citations are relative to /root/reference.
#pragma once
#include <cstdint>
#include <cstddef>
#include <vector>

namespace oracle {

// Layout-compatible with cv::KeyPoint (28 bytes): pt.x pt.y size angle response octave class_id
struct KeyPoint {
    float x, y, size, angle, response;
    int octave, class_id;
};

enum BlurVariant { BLUR_CV4_BITEXACT = 0, BLUR_LEGACY_24 = 1 };
enum RetainVariant { RETAIN_NTH_AT_N = 0, RETAIN_NTH_AT_N_MINUS_1 = 1 };
enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

// ---------------------------------------------------------------- OpenCV primitive restatements
int cvRound(double v);
int cvRoundf(float v);
float fastAtan2(float y, float x);
// 8UC1 bilinear resize (cv::resize INTER_LINEAR)
void resizeLinear8u(const uint8_t* src, int sw, int sh, int sstep, uint8_t* dst, int dw, int dh, int dstep);
// copyMakeBorder(src -> dst, b,b,b,b, BORDER_REFLECT_101); dst is (w+2b) x (h+2b)
void copyMakeBorder101(const uint8_t* src, int w, int h, int sstep, uint8_t* dst, int dstep, int b);
// GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) on an ROI of a bordered plane, in place.
// `roi` points at ROI (0,0); pixels outside the ROI (>=3 px ring) must hold the reflect-101 border.
void gaussianBlur7x7InPlace(uint8_t* roi, int w, int h, int step, int variant);
// Isolated blur with computed reflect-101 border (used to validate the in-place claim).
void gaussianBlur7x7(const uint8_t* src, int w, int h, int sstep, uint8_t* dst, int dstep, int variant);
// cv::FAST(img, kps, threshold, nonmax=true), TYPE_9_16
extern bool g_fast_scalar_only;
void FAST9_16(const uint8_t* img, int w, int h, int step, int threshold, bool nonmax, std::vector<KeyPoint>& out);
// cv::KeyPointsFilter::retainBest
void retainBest(std::vector<KeyPoint>& kps, int n, int variant);

// ---------------------------------------------------------------- ORBextractor
class ORBextractor {
public:
    ORBextractor(int nfeatures = 1000, float scaleFactor = 1.2f, int nlevels = 8, int scoreType = FAST_SCORE,
                 int fastTh = 20, int minFastTh = 7, int blurVariant = BLUR_CV4_BITEXACT,
                 int retainVariant = RETAIN_NTH_AT_N);
    // returns 0 on success, <0 on a condition where the reference would assert/throw
    int operator()(const uint8_t* image, int w, int h, int step, std::vector<KeyPoint>& keypoints,
                   std::vector<uint8_t>& descriptors);
    int GetLevels() const { return nlevels; }
    float GetScaleFactor() const { return (float)scaleFactor; }

    // exposed for tests
    struct Level { int w, h, step; std::vector<uint8_t> buf; uint8_t* roi() { return buf.data() + 16 * step + 16; } };
    std::vector<Level> mvImagePyramid;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<int> umax;
    std::vector<float> mvScaleFactor, mvInvScaleFactor;
    std::vector<std::vector<KeyPoint>> lastAllKeypoints;  // per level, level coordinates, with angle
    void ComputePyramid(const uint8_t* image, int w, int h, int step);
    int ComputeKeyPoints(std::vector<std::vector<KeyPoint>>& all);
    int ComputeKeyPointsOctTree(std::vector<std::vector<KeyPoint>>& all);   // orb_quadtree.cpp: ORB-SLAM2's quadtree mode (non-default)
    int distribution = 0;                                                    // 0: the reference's cell quota, 1: quadtree

    int nfeatures;
    double scaleFactor;
    int nlevels, scoreType, fastTh, minFastTh, blurVariant, retainVariant;
};

// ---------------------------------------------------------------- Frame grid + ORBmatcher
static const int FRAME_GRID_COLS = 64, FRAME_GRID_ROWS = 48;

struct Frame {
    int N = 0;
    std::vector<KeyPoint> mvKeysUn;
    std::vector<uint8_t> mDescriptors;     // N x 32
    std::vector<int> mvpMapPoints;         // map-point id or -1 (NULL)
    std::vector<int> mvpMatchScore;        // 999 default
    std::vector<uint8_t> mvbOutlier;
    std::vector<float> mvScaleFactors, mvLevelSigma2;
    int mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0;
    float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
    float fx = 0, fy = 0, cx = 0, cy = 0;
    float mTcw[16] = {1,0,0,0, 0,1,0,0, 0,0,1,0, 0,0,0,1};
    std::vector<size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];

    void init(const KeyPoint* kps, const uint8_t* desc, int n, int imgW, int imgH, int nlevels, float scaleFactor);
    bool PosInGrid(const KeyPoint& kp, int& posX, int& posY) const;
    std::vector<size_t> GetFeaturesInArea(float x, float y, float r, int minLevel = -1, int maxLevel = -1) const;
};

// ---- cv::Mat (CV_32F) arithmetic of the matchers, restated from OpenCV's matop.cpp / matmul.cpp (2.4.13, 3.4.1 and 4.13 agree)
// y = R*x + t (3x3 * 3x1 + 3x1): MatOp_GEMM::add folds the sum into one cv::gemm(R, x, 1, t, 1); with no transpose flag and
// len = 3, cv::gemm takes its small-matrix path ("case 3" of the 2 <= len <= 4 switch): every row's dot product is accumulated
// in *float*, left to right, and d = (float)(t0*alpha + c*beta) is evaluated in double.  Pinned against cv2 4.13
// (tests/golden/cv413_gemm_small.npz).  rstep / tstep: element strides of the rows of R and of the entries of t.
inline void cvGemmRxPlusT(const float* R, int rstep, const float* x, const float* t, int tstep, float* y) {
    for (int r = 0; r < 3; r++) {
        const float* a = R + r * rstep;
        const float t0 = a[0] * x[0] + a[1] * x[1] + a[2] * x[2];
        y[r] = (float)((double)t0 * 1.0 + (double)t[r * tstep] * 1.0);
    }
}
// Ow = -R.t()*t: operator-(MatExpr) materialises the transpose (MatOp::subtract(Scalar, expr)) and scales it by -1; MatOp::matmul
// turns the scaled operand into cv::gemm(Rt, t, -1) with no transpose flag -> the same small-matrix path, c = 0
inline void cvNegRtTimesT(const float* R, int rstep, const float* t, int tstep, float* Ow) {
    for (int i = 0; i < 3; i++) {
        const float t0 = R[0 * rstep + i] * t[0] + R[1 * rstep + i] * t[tstep] + R[2 * rstep + i] * t[2 * tstep];
        Ow[i] = (float)((double)t0 * -1.0 + 0.0);
    }
}
// dst = src * alpha for a MatExpr scale (Mat / double, double * Mat): Mat::convertTo(CV_32F, alpha) multiplies by (float)alpha
inline float cvScale32f(float v, double alpha) { return v * (float)alpha + 0.0f; }

struct MapPointGeom {   // the MapPoint fields the mapping / loop-closing matchers read (include/MapPoint.h)
    float world[3], normal[3];       // GetWorldPos(), GetNormal()
    float minDist, maxDist;          // GetMinDistanceInvariance(), GetMaxDistanceInvariance()
    uint8_t desc[32];                // GetDescriptor()
};
struct Sim3Rt { float R[9], t[3], Ow[3]; };
// Scw (4x4 row-major) -> Rcw = sRcw/scw, tcw = Scw.col(3)/scw, Ow = -Rcw.t()*tcw (src/ORBmatcher.cc:866-871, :1718-1723)
Sim3Rt decomposeSim3(const float* Scw);

struct MapPointView {   // the MapPoint fields SearchByProjection reads (include/MapPoint.h)
    float mTrackProjX, mTrackProjY;
    int mnTrackScaleLevel;
    float mTrackViewCos;
    uint8_t mbTrackInView, isBad;
    uint8_t desc[32];
};

// single-keypoint entry points for known-answer tests (IC_Angle :131-158, computeOrbDescriptor :161-201)
float ic_angle_at(const uint8_t* img, int step, float x, float y);
void describe_at(const uint8_t* img, int step, float x, float y, float angleDeg, uint8_t* desc);

// Frame::isInFrustum (src/Frame.cc:166-227) on the fields it reads: Rcw (3x3 row-major), tcw, Ow, intrinsics, image
// bounds and mvScaleFactors.  Returns mbTrackInView and, when true, fills u, v, level, viewCos.
bool isInFrustum(const float* Rcw, const float* tcw, const float* Ow, float fx, float fy, float cx, float cy, int mnMinX,
                 int mnMaxX, int mnMinY, int mnMaxY, const std::vector<float>& mvScaleFactors, const float* P,
                 const float* Pn, float minDistance, float maxDistance, float viewingCosLimit, float& u, float& v,
                 int& level, float& viewCos);
// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:196-262): index of the representative descriptor among N
int distinctiveDescriptor(const uint8_t* descriptors, int N);
// Frame::UndistortKeyPoints' cv::undistortPoints call (src/Frame.cc:389-423), src / dst: n x 2 floats
void undistortPoints(const float* src, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist, float* dst);

class ORBmatcher {
public:
    static const int TH_LOW = 50, TH_HIGH = 100, HISTO_LENGTH = 30;
    ORBmatcher(float nnratio = 0.6f, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}
    static int DescriptorDistance(const uint8_t* a, const uint8_t* b);
    int SearchByProjection(Frame& F, const std::vector<MapPointView>& vpMapPoints, float th);
    int SearchByProjection_OnePoint(Frame& F, const MapPointView& mp, int mpId, float th);
    // good[i] = (goodAtFrameId == F.mnId) of map point i / of LastFrame.mvpMapPoints[i]
    int SearchByProjection_GoodFeature(Frame& F, const std::vector<MapPointView>& vpMapPoints, const uint8_t* good, float th_good,
                                       float th_rest);
    int SearchByProjection_GoodFeature(Frame& Cur, const Frame& Last, const float* lastWorldPos, const uint8_t* good,
                                       float th_good, float th_rest);
    int SearchByProjection_Budget(Frame& F, const std::vector<MapPointView>& vpMapPoints, float th, double time_constr);
    // last-frame variant: world positions of LastFrame's map points are given per keypoint (or id<0)
    int SearchByProjection(Frame& Cur, const Frame& Last, const float* lastWorldPos /*N x 3*/, float th);
    int SearchForInitialization(const Frame& F1, const Frame& F2, std::vector<float>& vbPrevMatched /*N1 x 2*/,
                                std::vector<int>& vnMatches12, int windowSize);
    int SearchByProjection(const Frame& F1, Frame& F2, const float* f1World /*N1 x 3*/, int windowSize, std::vector<int>& vpMapPointMatches2);
    // :724-853 (TrackReferenceKeyFrame / Relocalization) and :2204-2336 (Relocalization)
    int SearchByBoW(const Frame& KF, const int* kfNode, const Frame& F, const int* fNode, std::vector<int>& vpMapPointMatches);
    int SearchByProjection(Frame& CurrentFrame, const Frame& KF, const uint8_t* kfSkip, const float* kfWorld, const float* kfMinDist,
                           const uint8_t* kfMpDesc, float th, int ORBdist);
    // orb_matcher_kf.cpp: the overloads of the mapping / loop-closing threads.  Map points are ids (>= 0) in
    // Frame::mvpMapPoints (-1 = NULL); `bad` / geometry arrays are indexed by id.
    int SearchByBoW_KeyFrames(const Frame& KF1, const int* node1, const Frame& KF2, const int* node2, std::vector<int>& vpMatches12);   // SearchByBoW(pKF1, pKF2, ..) :1289-1424
    int SearchForTriangulation(const Frame& KF1, const int* node1, const Frame& KF2, const int* node2, const float* F12 /*3x3*/,
                               std::vector<int>& vMatches12);                                                                    // :1426-1588
    // Fuse: `points` = vpMapPoints as ids (-1 = NULL); state (in/out): kfSlot = pKF->mvpMapPoints as ids, bad[id], idxInKF[id]
    // (-1: pKF does not observe the point).  bestIdx / bestDist per list position (-1 / INT_MAX where the point was skipped).
    int Fuse(const Frame& KF, const std::vector<int>& points, const MapPointGeom* geom, float th, std::vector<int>& kfSlot,
             std::vector<uint8_t>& bad, std::vector<int>& idxInKF, std::vector<int>& bestIdx, std::vector<int>& bestDist);       // :1590-1708
    int Fuse(const Frame& KF, const float* Scw, const std::vector<int>& points, const MapPointGeom* geom, float th,
             std::vector<int>& kfSlot, std::vector<uint8_t>& bad, std::vector<int>& idxInKF, std::vector<int>& bestIdx,
             std::vector<int>& bestDist);                                                                                        // :1710-1839
    int SearchByProjection(const Frame& KF, const float* Scw, const std::vector<int>& points, const MapPointGeom* geom,
                           const uint8_t* bad, std::vector<int>& vpMatched, int th);                                             // :855-977
    // vpMatches12 (in/out, per KF1 keypoint): in: id or -1; out: newly found pairs hold vpMapPoints2[idx2].  idx2InKF2[i1] =
    // pMP->GetIndexInKeyFrame(pKF2) of vpMatches12[i1].  geom / bad are indexed by map-point id.
    int SearchBySim3(const Frame& KF1, const Frame& KF2, std::vector<int>& vpMatches12, const int* idx2InKF2, const MapPointGeom* geom,
                     const uint8_t* bad, float s12, const float* R12, const float* t12, float th);                               // :1841-2079
    static bool CheckDistEpipolarLine(const KeyPoint& kp1, const KeyPoint& kp2, const float* F12, float sigma2kp2);              // :705-722
    int WindowSearch(const Frame& F1, const Frame& F2, int windowSize, std::vector<int>& vpMapPointMatches2,
                     int minScaleLevel = -1, int maxScaleLevel = 0x7fffffff);
    static void ComputeThreeMaxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3);
    static float RadiusByViewingCos(float viewCos);
    float mfNNratio;
    bool mbCheckOrientation;
};

// ---------------------------------------------------------------- Observability (good features)
struct PinHoleCamera {
    double fu, fv; int nRows, nCols; double Cx, Cy, k1, k2, dx, dy;
};

struct KineStruct {
    double dt, dt_inSeg;
    double Xv[13];
    double F_Q[16], F_Omg[12], F[169];              // row-major
    double F_Q_inSeg[16], F_Omg_inSeg[12], F_inSeg[169];
    float Tcw[16];
};

struct GoodPoint { size_t idx; double obs_score; double obs_block[49]; bool selected; double upper_bound; };

class Observability {
public:
    Observability(double f, int nRows, int nCols, double Cx, double Cy, double k1, double k2, double dx, double dy);
    Observability(double fu, double fv, int nRows, int nCols, double Cx, double Cy, double k1, double k2);
    void updatePWLSVec(double time_prev, const float* Tcw_prev, double time_cur, const float* Twc_cur);
    void predictPWLSVec(double dt, size_t num_seg_pred);

    // per-point Jacobians; all matrices row-major
    bool compute_H_subblock_simplied(const double* Xv, const double* yi, double* H13 /*2x3*/, double* H47 /*2x4*/,
                                     double* dhu_dhrl /*2x3*/, bool check_viz, float& u, float& v) const;
    void compute_H_subblock_complete(const double* Xv, const double* yi, const double* zi, double* H13, double* H47,
                                     double* dhu_dhrl, float& residual_u, float& residual_v) const;
    // batchInfoMat_Frame for every index (FRAME_INFO_MATRIX). sigma2 = scale[octave]^2 (float). Writes 49 doubles
    // per point into blocks (row-major), obsScore (-1 if not valid).
    void batchInfoMat_Frame(const float* worldPos, const int* octave, const int* matchScore, const uint8_t* valid,
                            int n, const float* levelSigma2, double* blocks, double* obsScore, double* H_meas /*n x 14 or null*/);
    void batchInfoMat_Map(const float* worldPos, const uint8_t* trackInView, int n, bool check_viz,
                          int mnMinX, int mnMaxX, int mnMinY, int mnMaxY, double* blocks, double* obsScore,
                          float* uv /*n x 2 or null*/, double* H_meas = nullptr /*n x 14 or null*/);

    // subset selection over lmkSelectPool
    std::vector<GoodPoint> lmkSelectPool;
    void maxVolSelection_BaselineGreedy(std::vector<GoodPoint>* subVec, size_t mpLimit);
    bool maxVolSelection_LazierGreedy(size_t stIdx, size_t edIdx, std::vector<GoodPoint>* subVec, size_t mpLimit,
                                      double sampleScale, double max_time_for_select);
    bool maxVolDeletion_LazierGreedy(size_t stIdx, size_t edIdx, std::vector<GoodPoint>* subVec, size_t mpLimit,
                                     double sampleScale, double max_time_for_select);
    bool maxVolAutomatic_LazierGreedy(size_t stIdx, size_t edIdx, std::vector<GoodPoint>* subVec, size_t mpLimit,
                                      double sampleScale, double max_time_for_select);

    PinHoleCamera camera;
    double Xv[13];
    std::vector<KineStruct> kinematic;
    size_t mKineIdx = 0;
    double mBoundXInFrame = 20, mBoundYInFrame = 20, mBoundDepth = 0;
    int vizMinX = 0, vizMaxX = 0, vizMinY = 0, vizMaxY = 0;   // Frame::mnMinX.. statics (include/Frame.h:142-145)
    // RNG hook: the reference uses the process-global glibc rand(); tests may swap in a recorded stream.
    int (*rand_fn)() = nullptr;
};

double logDet(const double* M, int n);   // include/Util.hpp:714-731

// Observability::runActiveMapMatching (src/Observability.cc:1249-1524) for FRAME_INFO_MATRIX / MAP_INFO_MATRIX with
// INFORMATION_EFFICIENCY_SCORE off and an infinite time cap.  inPool[i] = (updateAtFrameId == pFrame->mnId),
// obsMat = MapPoint::ObsMat (M x 49), H_meas = MapPoint::H_meas (M x 14, rows [H13 H47]); curMat (49, in/out) enters as
// mBaseInfoMat and leaves as the accumulated matrix; left = mLeftMapPoints (indices into mps, appended).
int runActiveMapMatching(Frame& F, const std::vector<MapPointView>& mps, const uint8_t* inPool, const double* obsMat,
                         const double* H_meas, double* curMat, float th, ORBmatcher& matcher, int num_to_match,
                         std::vector<int>& left, int (*rand_fn)() = nullptr);

// Optimizer::PoseOptimization (src/Optimizer.cc:279-413), pose_optimizer.cpp
int PoseOptimization(float* Tcw, int N, const float* kpXY, const int* octave, const int* mapPoint, const float* world,
                     const float* invLevelSigma2, float fx, float fy, float cx, float cy, uint8_t* outlier);

}  // namespace oracle
