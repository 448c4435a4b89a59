// orb_slam.hpp -- header-only C++ mirror of the reference's operator API over the C ABI in gforb.h.
//
// A GF-ORB-SLAM maintainer replaces
//     #include "ORBextractor.h" / "ORBmatcher.h" / "Observability.h"
// by this header (and links libgforb.so) to run the per-frame tracking front end on a B200.  Class and member names,
// argument order and meaning, defaults and error behaviour follow the reference:
//     ORB_SLAM::ORBextractor   include/ORBextractor.h:51-104,  src/ORBextractor.cc:464-998
//     ORB_SLAM::ORBmatcher     include/ORBmatcher.h:39-294,    src/ORBmatcher.cc
//     ORB_SLAM::Observability  include/Observability.h:113-770, src/Observability.cc
// With -DGFORB_WITH_OPENCV the real cv::Mat / cv::KeyPoint / cv::Point2f are used; otherwise layout-compatible
// stand-ins (namespace gforb_cv) keep the header free of dependencies.  Frame and MapPoint are reduced to the fields
// this path reads and writes (the shim flattens them into the SoA buffers of the C ABI and scatters results back).
// There is no CPU fallback: if no CUDA device is usable the constructors throw std::runtime_error.
#ifndef GFORB_ORB_SLAM_HPP
#define GFORB_ORB_SLAM_HPP

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <map>
#include <set>
#include <vector>

#include "../gforb.h"

#ifdef GFORB_WITH_OPENCV
#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>
namespace gforb_cv = cv;
#else
namespace gforb_cv {
struct Point2f { float x = 0, y = 0; Point2f() {} Point2f(float x_, float y_) : x(x_), y(y_) {} };
struct KeyPoint {            // same 28-byte layout as cv::KeyPoint
    Point2f pt; float size = 0, angle = -1, response = 0; int octave = 0, class_id = -1;
};
// 8-bit single-channel matrix (CV_8UC1) or, for poses, 32-bit float (CV_32F); owns its storage unless wrapping
struct Mat {
    int rows = 0, cols = 0, elem = 1;
    size_t step = 0;
    uint8_t* data = nullptr;
    std::vector<uint8_t> own;
    Mat() {}
    Mat(const Mat& o) { *this = o; }
    Mat& operator=(const Mat& o) {       // an owning matrix re-binds its data pointer to the copied storage
        rows = o.rows; cols = o.cols; elem = o.elem; step = o.step; own = o.own;
        data = (!o.own.empty() && o.data == o.own.data()) ? own.data() : o.data;
        return *this;
    }
    Mat(int r, int c, uint8_t* ext, size_t stp) : rows(r), cols(c), elem(1), step(stp), data(ext) {}
    bool empty() const { return rows == 0 || cols == 0 || data == nullptr; }
    void create(int r, int c, int elemBytes = 1) {
        rows = r; cols = c; elem = elemBytes; step = (size_t)c * elemBytes;
        own.assign((size_t)r * step, 0); data = own.data();
    }
    uint8_t* ptr(int r = 0) { return data + (size_t)r * step; }
    const uint8_t* ptr(int r = 0) const { return data + (size_t)r * step; }
    Mat row(int r) const { return Mat(1, cols, const_cast<uint8_t*>(ptr(r)), step); }
    template <class T> T& at(int r, int c) { return reinterpret_cast<T*>(ptr(r))[c]; }
    template <class T> const T& at(int r, int c) const { return reinterpret_cast<const T*>(ptr(r))[c]; }
};
}  // namespace gforb_cv
#endif

namespace ORB_SLAM {

static_assert(sizeof(gforb_cv::KeyPoint) == sizeof(GforbKeyPoint), "KeyPoint must be layout-compatible with cv::KeyPoint");

inline void gforb_throw(int rc, const char* what) {
    if (rc == GFORB_OK) return;
    throw std::runtime_error(std::string(what) + ": gforb error " + std::to_string(rc));
}

// ------------------------------------------------------------------------------------------------ ORBextractor
class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };
    ORBextractor(int nfeatures = 1000, float scaleFactor = 1.2f, int nlevels = 8, int scoreType = FAST_SCORE,
                 int fastTh = 20, int device = 0)
        : nfeatures_(nfeatures), scaleFactor_(scaleFactor), nlevels_(nlevels), scoreType_(scoreType), fastTh_(fastTh),
          device_(device) {
        if (gforb_device_count() <= 0) throw std::runtime_error("ORBextractor: no CUDA device (the B200 path has no CPU fallback)");
    }
    ~ORBextractor() { if (h_) gforb_extractor_destroy(h_); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // Compute the ORB features and descriptors on an image (include/ORBextractor.h:60-64).  The mask is ignored, as in
    // the reference (always empty, src/Frame.cc:59).  An empty image returns silently (src/ORBextractor.cc:772-773).
    void operator()(const gforb_cv::Mat& image, const gforb_cv::Mat& /*mask*/, std::vector<gforb_cv::KeyPoint>& keypoints,
                    gforb_cv::Mat& descriptors) {
        if (image.empty()) return;
        ensure(image.cols, image.rows);
        kp_.resize(nfeatures_);
        desc_.resize((size_t)nfeatures_ * 32);
        int32_t n = 0;
        gforb_throw(gforb_extract(h_, image.data, image.cols, image.rows, (int)image.step, image.step * image.rows, 1,
                                  kp_.data(), desc_.data(), &n, 0, nullptr), "ORBextractor::operator()");
        keypoints.clear();                                   // src/ORBextractor.cc:816
        if (n == 0) { descriptors = gforb_cv::Mat(); return; }
        keypoints.resize(n);
        std::memcpy(static_cast<void*>(keypoints.data()), kp_.data(), (size_t)n * sizeof(GforbKeyPoint));
#ifdef GFORB_WITH_OPENCV
        descriptors.create(n, 32, CV_8U);
#else
        descriptors.create(n, 32);
#endif
        for (int i = 0; i < n; i++) std::memcpy(descriptors.ptr(i), desc_.data() + (size_t)i * 32, 32);
    }
    int GetLevels() { return nlevels_; }
    float GetScaleFactor() { return scaleFactor_; }
    // Not in the reference: 0 = its cell-quota keypoint distribution (default), 1 = ORB-SLAM2's quadtree distribution
    // (gforb_extractor_set_distribution); applies to the next operator() call.
    void SetKeypointDistribution(int mode) {
        distribution_ = mode;
        if (h_) gforb_throw(gforb_extractor_set_distribution(h_, mode), "ORBextractor::SetKeypointDistribution");
    }

private:
    void ensure(int w, int h) {
        if (h_ && w <= maxW_ && h <= maxH_) return;
        if (h_) { gforb_extractor_destroy(h_); h_ = nullptr; }
        GforbExtractorConfig c;
        std::memset(&c, 0, sizeof(c));
        c.nfeatures = nfeatures_; c.scaleFactor = scaleFactor_; c.nlevels = nlevels_; c.scoreType = scoreType_;
        c.iniThFAST = fastTh_; c.minThFAST = 7; c.blur_variant = 0; c.retain_variant = 0;
        c.max_width = w; c.max_height = h; c.max_batch = 1; c.device = device_;
        gforb_throw(gforb_extractor_create(&c, &h_), "ORBextractor: create");
        if (distribution_) gforb_throw(gforb_extractor_set_distribution(h_, distribution_), "ORBextractor: distribution");
        maxW_ = w; maxH_ = h;
    }
    int nfeatures_; float scaleFactor_; int nlevels_, scoreType_, fastTh_, device_;
    GforbExtractor* h_ = nullptr;
    int maxW_ = 0, maxH_ = 0, distribution_ = 0;
    std::vector<GforbKeyPoint> kp_;
    std::vector<uint8_t> desc_;
};

// ------------------------------------------------------------------------------------------------ Frame / MapPoint views
// The MapPoint fields the tracking matchers read (include/MapPoint.h): projection cache written by Frame::isInFrustum,
// the representative descriptor and the world position.
struct KeyFrame;
struct MapPoint {
    float mTrackProjX = 0, mTrackProjY = 0;
    int mnTrackScaleLevel = 0;
    float mTrackViewCos = 0;
    bool mbTrackInView = false, mbBad = false;
    uint8_t mDescriptor[32] = {0};
    float mWorldPos[3] = {0, 0, 0};
    float mNormalVector[3] = {0, 0, 1};              // MapPoint::GetNormal()
    float mfMinDistance = 0, mfMaxDistance = 0;      // GetMinDistanceInvariance() / GetMaxDistanceInvariance()
    double ObsScore = -1;
    // written by Observability::runMatrixBuilding(MAP_INFO_MATRIX) (src/Observability.cc:627-636)
    float u_proj = 0, v_proj = 0;
    double H_meas[14] = {0};                         // 2x7 row-major [H13 H47]
    double ObsMat[49] = {0};                         // 7x7 row-major
    long unsigned int updateAtFrameId = 0;
    long unsigned int goodAtFrameId = 0;             // good-feature stamp read by SearchByProjection_GoodFeature
    bool isBad() const { return mbBad; }
    const uint8_t* GetDescriptor() const { return mDescriptor; }
    // observations (src/MapPoint.cc:77-81, :270-283, :136-170): what the Fuse overloads and SearchBySim3 read and write
    std::map<KeyFrame*, size_t> mObservations;
    void AddObservation(KeyFrame* pKF, size_t idx) { mObservations[pKF] = idx; }
    bool IsInKeyFrame(KeyFrame* pKF) const { return mObservations.count(pKF) != 0; }
    int GetIndexInKeyFrame(KeyFrame* pKF) const {
        auto it = mObservations.find(pKF);
        return it == mObservations.end() ? -1 : (int)it->second;
    }
    inline void Replace(MapPoint* pMP);                // defined after KeyFrame
};

// The Frame fields this path touches (include/Frame.h): undistorted keypoints, descriptors, per-keypoint map-point
// association / match score / outlier flag, image bounds and the scale pyramid parameters.
struct Frame {
    long unsigned int mnId = 0;
    int N = 0;
    std::vector<gforb_cv::KeyPoint> mvKeysUn;
    gforb_cv::Mat mDescriptors;                      // N x 32
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<int> mvpMatchScore;
    std::vector<bool> mvbOutlier, mvbCandidate, mvbGoodFeature;
    int mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0;   // statics in the reference (include/Frame.h:142-145)
    int mnScaleLevels = 8;
    float mfScaleFactor = 1.2f;
    std::vector<float> mvLevelSigma2;
    float mTcw[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    float mRcw[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, mtcw[3] = {0, 0, 0}, mOw[3] = {0, 0, 0};   // Frame::UpdatePoseMatrices
    float fx = 0, fy = 0, cx = 0, cy = 0;
    std::map<unsigned, std::vector<unsigned>> mFeatVec;   // DBoW2::FeatureVector (node id -> feature indices), filled by the caller's ComputeBoW

    Frame() {}
    // what Frame::Frame does for this path after extraction (src/Frame.cc:40-131; no distortion -> mvKeysUn = mvKeys)
    Frame(const std::vector<gforb_cv::KeyPoint>& keys, const gforb_cv::Mat& desc, int imgW, int imgH, int nlevels = 8,
          float scaleFactor = 1.2f)
        : N((int)keys.size()), mvKeysUn(keys), mDescriptors(desc), mvpMapPoints(keys.size(), nullptr),
          mvpMatchScore(keys.size(), 999), mvbOutlier(keys.size(), false), mvbCandidate(keys.size(), true),
          mvbGoodFeature(keys.size(), false), mnMinX(0), mnMaxX(imgW), mnMinY(0), mnMaxY(imgH), mnScaleLevels(nlevels),
          mfScaleFactor(scaleFactor) {
        mvLevelSigma2.assign(nlevels, 1.0f);
        float s = 1.0f;
        for (int i = 1; i < nlevels; i++) { s = s * scaleFactor; mvLevelSigma2[i] = s * s; }
    }
};

// The KeyFrame fields the Tracking-side matcher overloads read (include/KeyFrame.h): GetMapPointMatches(), GetFeatureVector(),
// GetDescriptor(i), GetKeyPointUn(i).  Built from a Frame like the reference's KeyFrame::KeyFrame(Frame&, Map*, KeyFrameDatabase*).
struct KeyFrame {
    int N = 0;
    std::vector<gforb_cv::KeyPoint> mvKeysUn;
    gforb_cv::Mat mDescriptors;
    std::vector<MapPoint*> mvpMapPoints;
    std::map<unsigned, std::vector<unsigned>> mFeatVec;
    int mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0;
    int mnScaleLevels = 8;
    float mfScaleFactor = 1.2f;
    float Tcw[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};   // GetPose(): GetRotation() / GetTranslation() are its blocks
    float fx = 0, fy = 0, cx = 0, cy = 0;
    KeyFrame() {}
    explicit KeyFrame(const Frame& F)
        : N(F.N), mvKeysUn(F.mvKeysUn), mDescriptors(F.mDescriptors), mvpMapPoints(F.mvpMapPoints), mFeatVec(F.mFeatVec),
          mnMinX(F.mnMinX), mnMaxX(F.mnMaxX), mnMinY(F.mnMinY), mnMaxY(F.mnMaxY), mnScaleLevels(F.mnScaleLevels),
          mfScaleFactor(F.mfScaleFactor), fx(F.fx), fy(F.fy), cx(F.cx), cy(F.cy) { std::memcpy(Tcw, F.mTcw, sizeof(Tcw)); }
    std::vector<MapPoint*> GetMapPointMatches() const { return mvpMapPoints; }
    const std::map<unsigned, std::vector<unsigned>>& GetFeatureVector() const { return mFeatVec; }
    gforb_cv::KeyPoint GetKeyPointUn(size_t i) const { return mvKeysUn[i]; }
    // src/KeyFrame.cc:205-251
    void AddMapPoint(MapPoint* pMP, size_t idx) { mvpMapPoints[idx] = pMP; }
    void EraseMapPointMatch(size_t idx) { mvpMapPoints[idx] = nullptr; }
    void ReplaceMapPointMatch(size_t idx, MapPoint* pMP) { mvpMapPoints[idx] = pMP; }
    MapPoint* GetMapPoint(size_t idx) const { return mvpMapPoints[idx]; }
    std::set<MapPoint*> GetMapPoints() const {
        std::set<MapPoint*> s;
        for (MapPoint* p : mvpMapPoints) if (p && !p->isBad()) s.insert(p);
        return s;
    }
};

// src/MapPoint.cc:136-170 without the map bookkeeping that is not on this path (ComputeDistinctiveDescriptors is
// ORBmatcher::ComputeDistinctiveDescriptors below; Map::EraseMapPoint is the caller's)
inline void MapPoint::Replace(MapPoint* pMP) {
    if (pMP == this) return;
    std::map<KeyFrame*, size_t> obs = mObservations;
    mObservations.clear();
    mbBad = true;
    for (auto& o : obs) {
        KeyFrame* pKF = o.first;
        if (!pMP->IsInKeyFrame(pKF)) {
            pKF->ReplaceMapPointMatch(o.second, pMP);
            pMP->AddObservation(pKF, o.second);
        } else {
            pKF->EraseMapPointMatch(o.second);
        }
    }
}

// ------------------------------------------------------------------------------------------------ ORBmatcher
class ORBmatcher {
public:
    static const int TH_LOW = 50, TH_HIGH = 100, HISTO_LENGTH = 30;      // src/ORBmatcher.cc:40-42
    ORBmatcher(float nnratio = 0.6f, bool checkOri = true, int device = 0)
        : mfNNratio(nnratio), mbCheckOrientation(checkOri), device_(device) {
        if (gforb_device_count() <= 0) throw std::runtime_error("ORBmatcher: no CUDA device (the B200 path has no CPU fallback)");
    }
    ~ORBmatcher() { if (h_) gforb_matcher_destroy(h_); }
    ORBmatcher(const ORBmatcher&) = delete;
    ORBmatcher& operator=(const ORBmatcher&) = delete;

    // Hamming distance of two 32-byte descriptors (src/ORBmatcher.cc:2384-2400).  The reference calls this scalar helper
    // from host loops; here it goes through gforb_descriptor_distance like every batched caller (n = 1).
    // `static` like the reference (call sites write ORBmatcher::DescriptorDistance(a, b)): the call borrows a process-wide
    // matcher handle on device 0, created on first use.
    static int DescriptorDistance(const gforb_cv::Mat& a, const gforb_cv::Mat& b) {
        static ORBmatcher shared(0.6f, true, 0);
        shared.ensure(1, 1, 8, 1.2f);
        int32_t d = 0;
        gforb_throw(gforb_descriptor_distance(shared.h_, a.ptr(0), b.ptr(0), 1, &d, 0, nullptr), "DescriptorDistance");
        return d;
    }

    // Search matches between Frame keypoints and projected MapPoints; returns the number of matches
    // (include/ORBmatcher.h:48-50, src/ORBmatcher.cc:384-465).  F.mvpMapPoints / F.mvpMatchScore are updated in place.
    int SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th = 3) {
        return searchMap(F, vpMapPoints, th, 0.f, false);
    }
    // include/ORBmatcher.h:52-53, src/ORBmatcher.cc:51-140: window x th_good for points with goodAtFrameId == F.mnId,
    // x th_rest for the others; F.mvpMatchScore is left alone by this overload.
    int SearchByProjection_GoodFeature(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th_good, const float th_rest) {
        return searchMap(F, vpMapPoints, th_rest, th_good, true);
    }
    // include/ORBmatcher.h:67-68, src/ORBmatcher.cc:276-379: a spent budget matches nothing; this path has no wall-clock
    // cut-off inside the search, so a positive budget is the plain search.
    int SearchByProjection_Budget(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th, const double time_constr) {
        if (time_constr <= 0) return 0;
        return searchMap(F, vpMapPoints, th, 0.f, false);
    }

private:
    int searchMap(Frame& F, const std::vector<MapPoint*>& vpMapPoints, float th, float thGood, bool goodMode) {
        const int M = (int)vpMapPoints.size();
        if (F.N == 0 || M == 0) return 0;
        ensure(F.N, M, F.mnScaleLevels, F.mfScaleFactor);
        std::vector<float> u(M), v(M), vc(M);
        std::vector<int32_t> lvl(M);
        std::vector<uint8_t> fl(M), desc((size_t)M * 32);
        for (int i = 0; i < M; i++) {
            const MapPoint* p = vpMapPoints[i];
            u[i] = p->mTrackProjX; v[i] = p->mTrackProjY; vc[i] = p->mTrackViewCos; lvl[i] = p->mnTrackScaleLevel;
            fl[i] = (uint8_t)((p->mbTrackInView ? 1 : 0) | (p->isBad() ? 2 : 0) | ((goodMode && p->goodAtFrameId == F.mnId) ? 4 : 0));
            std::memcpy(&desc[(size_t)i * 32], p->GetDescriptor(), 32);
        }
        FlatFrame ff(F);
        std::vector<int32_t> ids = ids_of(F, vpMapPoints), score(F.mvpMatchScore.begin(), F.mvpMatchScore.end());
        int32_t mcount = M, n = 0;
        GforbMapPointSet mp = {u.data(), v.data(), lvl.data(), vc.data(), fl.data(), desc.data(), &mcount, M};
        if (goodMode)
            gforb_throw(gforb_search_by_projection_good_feature(h_, &ff.set, &mp, 1, thGood, th, mfNNratio, ids.data(), &n, 0, nullptr),
                        "SearchByProjection_GoodFeature");
        else
            gforb_throw(gforb_search_by_projection(h_, &ff.set, &mp, 1, th, mfNNratio, ids.data(), score.data(), &n, 0, nullptr),
                        "SearchByProjection");
        for (int i = 0; i < F.N; i++) {
            if (ids[i] >= 0 && ids[i] < M) F.mvpMapPoints[i] = vpMapPoints[ids[i]];
            if (!goodMode) F.mvpMatchScore[i] = score[i];
        }
        return n;
    }

public:

    // Frame::isInFrustum(pMP, viewingCosLimit) (src/Frame.cc:166-227) for every point of vpMapPoints in one call
    // (Tracking::SearchReferencePointsInFrustum loops over the local map, src/Tracking.cc): sets mbTrackInView and,
    // for the points in view, mTrackProjX / mTrackProjY / mnTrackScaleLevel / mTrackViewCos.  Returns how many are in view.
    int isInFrustum(Frame& F, std::vector<MapPoint*>& vpMapPoints, float viewingCosLimit) {
        const int M = (int)vpMapPoints.size();
        if (M == 0) return 0;
        ensure(std::max(F.N, 1), M, F.mnScaleLevels, F.mfScaleFactor);
        std::vector<float> world((size_t)M * 3), normal((size_t)M * 3), mind(M), maxd(M), u(M), v(M), vc(M);
        std::vector<int32_t> lvl(M);
        std::vector<uint8_t> fl(M, 0);
        for (int i = 0; i < M; i++) {
            const MapPoint* p = vpMapPoints[i];
            std::memcpy(&world[(size_t)i * 3], p->mWorldPos, 12); std::memcpy(&normal[(size_t)i * 3], p->mNormalVector, 12);
            mind[i] = p->mfMinDistance; maxd[i] = p->mfMaxDistance;
        }
        int32_t cnt = M;
        gforb_throw(gforb_frame_is_in_frustum(h_, 1, F.mRcw, F.mtcw, F.mOw, F.fx, F.fy, F.cx, F.cy, F.mnMinX, F.mnMaxX, F.mnMinY,
                                              F.mnMaxY, world.data(), normal.data(), mind.data(), maxd.data(), &cnt, M,
                                              viewingCosLimit, u.data(), v.data(), lvl.data(), vc.data(), fl.data(), 0, nullptr),
                    "isInFrustum");
        int n = 0;
        for (int i = 0; i < M; i++) {
            MapPoint* p = vpMapPoints[i];
            p->mbTrackInView = (fl[i] & 1) != 0;
            if (!p->mbTrackInView) continue;
            p->mTrackProjX = u[i]; p->mTrackProjY = v[i]; p->mnTrackScaleLevel = lvl[i]; p->mTrackViewCos = vc[i];
            n++;
        }
        return n;
    }

    // MapPoint::ComputeDistinctiveDescriptors() (src/MapPoint.cc:196-262) for a set of map points at once:
    // vObservationDescriptors[i] = the descriptors of point i's non-bad observations (32 bytes each, concatenated);
    // the representative descriptor is written to vpMapPoints[i]->mDescriptor (points without observations keep theirs).
    void ComputeDistinctiveDescriptors(const std::vector<MapPoint*>& vpMapPoints,
                                       const std::vector<std::vector<uint8_t>>& vObservationDescriptors) {
        const int n = (int)vpMapPoints.size();
        if (n == 0) return;
        ensure(1, 1, nlevels_ > 0 ? nlevels_ : 8, sf_ > 0 ? sf_ : 1.2f);
        std::vector<int32_t> off(n + 1, 0), best(n, -1);
        for (int i = 0; i < n; i++) off[i + 1] = off[i] + (int32_t)(vObservationDescriptors[i].size() / 32);
        std::vector<uint8_t> all((size_t)off[n] * 32), out((size_t)n * 32);
        for (int i = 0; i < n; i++)
            if (off[i + 1] > off[i]) std::memcpy(&all[(size_t)off[i] * 32], vObservationDescriptors[i].data(), (size_t)(off[i + 1] - off[i]) * 32);
        gforb_throw(gforb_mappoint_distinctive_descriptors(h_, n, all.data(), off.data(), best.data(), out.data(), 0, nullptr),
                    "ComputeDistinctiveDescriptors");
        for (int i = 0; i < n; i++)
            if (best[i] >= 0) std::memcpy(vpMapPoints[i]->mDescriptor, &out[(size_t)i * 32], 32);
    }

    // Frame::UndistortKeyPoints() (src/Frame.cc:389-423): mvKeysUn = mvKeys with cv::undistortPoints(pt, K, mDistCoef,
    // R = I, P = K); mDistCoef = (k1, k2, p1, p2[, k3]); mDistCoef[0] == 0 copies.
    void UndistortKeyPoints(const Frame& F, const std::vector<gforb_cv::KeyPoint>& mvKeys, const std::vector<float>& mDistCoef,
                            std::vector<gforb_cv::KeyPoint>& mvKeysUn) {
        const int n = (int)mvKeys.size();
        mvKeysUn.resize(mvKeys.size());
        if (n == 0) return;
        ensure(n, 1, F.mnScaleLevels, F.mfScaleFactor);
        static_assert(sizeof(gforb_cv::KeyPoint) == sizeof(GforbKeyPoint), "KeyPoint layout");
        int32_t cnt = n;
        gforb_throw(gforb_frame_undistort_keypoints(h_, 1, reinterpret_cast<const GforbKeyPoint*>(mvKeys.data()), &cnt, n, F.fx, F.fy,
                                                    F.cx, F.cy, mDistCoef.data(), (int)mDistCoef.size(),
                                                    reinterpret_cast<GforbKeyPoint*>(mvKeysUn.data()), 0, nullptr),
                    "UndistortKeyPoints");
    }

    // Matching for the map initialisation / frame-to-frame tracking without a map
    // (include/ORBmatcher.h:79-80, src/ORBmatcher.cc:1172-1287).
    int SearchForInitialization(Frame& F1, Frame& F2, std::vector<gforb_cv::Point2f>& vbPrevMatched,
                                std::vector<int>& vnMatches12, int windowSize = 10) {
        vnMatches12 = std::vector<int>(F1.N, -1);                  // :1175
        if (F1.N == 0 || F2.N == 0) return 0;
        ensure(std::max(F1.N, F2.N), F1.N, F2.mnScaleLevels, F2.mfScaleFactor);
        FlatFrame a(F1), b(F2);
        std::vector<float> prev((size_t)F1.N * 2);
        for (int i = 0; i < F1.N; i++) { prev[2 * i] = vbPrevMatched[i].x; prev[2 * i + 1] = vbPrevMatched[i].y; }
        std::vector<int32_t> m12(F1.N, -1);
        int32_t n = 0;
        gforb_throw(gforb_search_for_initialization(h_, &a.set, &b.set, 1, prev.data(), m12.data(), &n, windowSize,
                                                    mfNNratio, mbCheckOrientation ? 1 : 0, 0, nullptr),
                    "SearchForInitialization");
        for (int i = 0; i < F1.N; i++) {
            vnMatches12[i] = m12[i];
            vbPrevMatched[i] = gforb_cv::Point2f(prev[2 * i], prev[2 * i + 1]);     // :1282-1284
        }
        return n;
    }

    // Project MapPoints tracked in the last frame into the current frame and search matches
    // (include/ORBmatcher.h:52-54, src/ORBmatcher.cc:2081-2202).
    int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, float th) {
        return searchLast(CurrentFrame, LastFrame, th, 0.f, false);
    }
    // include/ORBmatcher.h:55-56, src/ORBmatcher.cc:143-274 (the call of Tracking::TrackWithMotionModel, src/Tracking.cc:1442)
    int SearchByProjection_GoodFeature(Frame& CurrentFrame, const Frame& LastFrame, const float th_good, const float th_rest) {
        return searchLast(CurrentFrame, LastFrame, th_rest, th_good, true);
    }

private:
    int searchLast(Frame& CurrentFrame, const Frame& LastFrame, float th, float thGood, bool goodMode) {
        if (CurrentFrame.N == 0 || LastFrame.N == 0) return 0;
        ensure(std::max(CurrentFrame.N, LastFrame.N), LastFrame.N, CurrentFrame.mnScaleLevels, CurrentFrame.mfScaleFactor);
        FlatFrame cur(CurrentFrame), last(LastFrame);
        const int NL = LastFrame.N;
        std::vector<int32_t> lastIds(NL, -1);
        std::vector<uint8_t> outl(NL, 0), good(NL, 0);
        std::vector<float> world((size_t)NL * 3, 0.f);
        for (int i = 0; i < NL; i++) {
            const MapPoint* p = LastFrame.mvpMapPoints[i];
            if (!p) continue;
            lastIds[i] = i; outl[i] = LastFrame.mvbOutlier[i] ? 1 : 0;
            good[i] = p->goodAtFrameId == CurrentFrame.mnId ? 1 : 0;
            std::memcpy(&world[(size_t)i * 3], p->mWorldPos, 12);
        }
        std::vector<int32_t> ids(CurrentFrame.N, -1), score(CurrentFrame.mvpMatchScore.begin(), CurrentFrame.mvpMatchScore.end());
        for (int i = 0; i < CurrentFrame.N; i++) if (CurrentFrame.mvpMapPoints[i]) ids[i] = NL;   // already claimed
        int32_t n = 0;
        if (goodMode)
            gforb_throw(gforb_search_by_projection_frames_good_feature(
                            h_, &cur.set, &last.set, 1, lastIds.data(), outl.data(), good.data(), world.data(), CurrentFrame.mTcw,
                            CurrentFrame.fx, CurrentFrame.fy, CurrentFrame.cx, CurrentFrame.cy, thGood, th, mfNNratio,
                            mbCheckOrientation ? 1 : 0, ids.data(), &n, 0, nullptr),
                        "SearchByProjection_GoodFeature(Current, Last)");
        else
            gforb_throw(gforb_search_by_projection_frames(h_, &cur.set, &last.set, 1, lastIds.data(), outl.data(), world.data(),
                                                          CurrentFrame.mTcw, CurrentFrame.fx, CurrentFrame.fy, CurrentFrame.cx,
                                                          CurrentFrame.cy, th, mfNNratio, mbCheckOrientation ? 1 : 0, ids.data(),
                                                          score.data(), &n, 0, nullptr),
                        "SearchByProjection(Current, Last)");
        for (int i = 0; i < CurrentFrame.N; i++) {
            if (ids[i] >= 0 && ids[i] < NL) CurrentFrame.mvpMapPoints[i] = LastFrame.mvpMapPoints[ids[i]];
            else if (ids[i] < 0) CurrentFrame.mvpMapPoints[i] = nullptr;
            if (!goodMode) CurrentFrame.mvpMatchScore[i] = score[i];
        }
        return n;
    }

public:

    // src/ORBmatcher.cc:979-1086
    int WindowSearch(Frame& F1, Frame& F2, int windowSize, std::vector<MapPoint*>& vpMapPointMatches2,
                     int minScaleLevel = -1, int maxScaleLevel = 0x7fffffff) {
        vpMapPointMatches2 = std::vector<MapPoint*>(F2.N, static_cast<MapPoint*>(nullptr));
        if (F1.N == 0 || F2.N == 0) return 0;
        ensure(std::max(F1.N, F2.N), F1.N, F2.mnScaleLevels, F2.mfScaleFactor);
        FlatFrame a(F1), b(F2);
        std::vector<int32_t> ids1(F1.N, -1), m2(F2.N, -1);
        for (int i = 0; i < F1.N; i++) if (F1.mvpMapPoints[i] && !F1.mvpMapPoints[i]->isBad()) ids1[i] = i;
        int32_t n = 0;
        gforb_throw(gforb_window_search(h_, &a.set, &b.set, 1, ids1.data(), windowSize, minScaleLevel, maxScaleLevel,
                                        mfNNratio, mbCheckOrientation ? 1 : 0, m2.data(), &n, 0, nullptr), "WindowSearch");
        for (int i = 0; i < F2.N; i++) if (m2[i] >= 0) vpMapPointMatches2[i] = F1.mvpMapPoints[m2[i]];
        return n;
    }

    // src/ORBmatcher.cc:1089-1168 (Tracking::TrackPreviousFrame): F1's map points that F2 does not hold yet, projected with
    // F2.mTcw into a window of windowSize pixels on the same level.  vpMapPointMatches2 starts as F2.mvpMapPoints;
    // F2.mvpMatchScore is updated.
    int SearchByProjection(Frame& F1, Frame& F2, int windowSize, std::vector<MapPoint*>& vpMapPointMatches2) {
        vpMapPointMatches2 = F2.mvpMapPoints;
        if (F1.N == 0 || F2.N == 0) return 0;
        ensure(std::max(F1.N, F2.N), F1.N, F2.mnScaleLevels, F2.mfScaleFactor);
        FlatFrame a(F1), b(F2);
        // MapPoint* identity -> ids: a point of F1 is its F1 index; a point only F2 holds gets F1.N + its F2 index
        std::vector<int32_t> ids1(F1.N, -1), m2(F2.N, -1), sc(F2.mvpMatchScore.begin(), F2.mvpMatchScore.end());
        std::vector<float> world((size_t)F1.N * 3, 0.f);
        for (int i = 0; i < F1.N; i++) {
            MapPoint* p = F1.mvpMapPoints[i];
            if (!p || p->isBad()) continue;
            ids1[i] = i;
            std::memcpy(&world[(size_t)i * 3], p->mWorldPos, 12);
        }
        for (int j = 0; j < F2.N; j++) {
            MapPoint* p = F2.mvpMapPoints[j];
            if (!p) continue;
            m2[j] = F1.N + j;
            for (int i = 0; i < F1.N; i++) if (F1.mvpMapPoints[i] == p) { m2[j] = i; break; }
        }
        int32_t n = 0;
        gforb_throw(gforb_search_by_projection_window(h_, &a.set, &b.set, 1, ids1.data(), world.data(), F2.mTcw, F2.fx, F2.fy, F2.cx,
                                                      F2.cy, windowSize, mfNNratio, m2.data(), sc.data(), &n, 0, nullptr),
                    "SearchByProjection(F1, F2, windowSize)");
        for (int j = 0; j < F2.N; j++) {
            if (m2[j] >= 0 && m2[j] < F1.N) vpMapPointMatches2[j] = F1.mvpMapPoints[m2[j]];
            F2.mvpMatchScore[j] = sc[j];
        }
        return n;
    }

    // src/ORBmatcher.cc:724-853 (Tracking::TrackReferenceKeyFrame, Relocalization): matches between the keyframe's map points
    // and the frame's keypoints over the features that share a vocabulary node.  The FeatureVectors are the caller's
    // (DBoW2 transform); a feature listed under several nodes keeps the last one, as DBoW2 never does that.
    int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches) {
        vpMapPointMatches = std::vector<MapPoint*>(F.mvpMapPoints.size(), static_cast<MapPoint*>(nullptr));
        if (!pKF || pKF->N == 0 || F.N == 0) return 0;
        ensure(std::max(pKF->N, F.N), pKF->N, F.mnScaleLevels, F.mfScaleFactor);
        FlatFrame a(*pKF), b(F);
        std::vector<int32_t> ids(pKF->N, -1), kn(pKF->N, -1), fn(F.N, -1), m(F.N, -1);
        for (int i = 0; i < pKF->N; i++) if (pKF->mvpMapPoints[i] && !pKF->mvpMapPoints[i]->isBad()) ids[i] = i;
        for (const auto& nd : pKF->mFeatVec) for (unsigned i : nd.second) if ((int)i < pKF->N) kn[i] = (int32_t)nd.first;
        for (const auto& nd : F.mFeatVec) for (unsigned i : nd.second) if ((int)i < F.N) fn[i] = (int32_t)nd.first;
        int32_t n = 0;
        gforb_throw(gforb_search_by_bow(h_, &a.set, ids.data(), kn.data(), &b.set, fn.data(), 1, mfNNratio, mbCheckOrientation ? 1 : 0,
                                        m.data(), &n, 0, nullptr), "SearchByBoW(KeyFrame, Frame)");
        for (int j = 0; j < F.N; j++) if (m[j] >= 0) vpMapPointMatches[j] = pKF->mvpMapPoints[m[j]];
        return n;
    }

    // src/ORBmatcher.cc:2204-2336 (Tracking::Relocalization): project the keyframe's map points that were not found yet with
    // the current pose; CurrentFrame.mvpMapPoints / mvpMatchScore are updated in place.
    int SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound, const float th, const int ORBdist) {
        if (!pKF || pKF->N == 0 || CurrentFrame.N == 0) return 0;
        ensure(std::max(pKF->N, CurrentFrame.N), pKF->N, CurrentFrame.mnScaleLevels, CurrentFrame.mfScaleFactor);
        FlatFrame cur(CurrentFrame), kf(*pKF);
        const int NK = pKF->N;
        std::vector<int32_t> ids(NK, -1);
        std::vector<uint8_t> skip(NK, 0), mpd((size_t)NK * 32, 0);
        std::vector<float> world((size_t)NK * 3, 0.f), mind(NK, 1.f);
        for (int i = 0; i < NK; i++) {
            MapPoint* p = pKF->mvpMapPoints[i];
            if (!p) continue;
            ids[i] = i;
            skip[i] = (p->isBad() || sAlreadyFound.count(p)) ? 1 : 0;
            std::memcpy(&world[(size_t)i * 3], p->mWorldPos, 12);
            mind[i] = p->mfMinDistance;
            std::memcpy(&mpd[(size_t)i * 32], p->mDescriptor, 32);
        }
        std::vector<int32_t> cids(CurrentFrame.N, -1), score(CurrentFrame.mvpMatchScore.begin(), CurrentFrame.mvpMatchScore.end());
        for (int j = 0; j < CurrentFrame.N; j++) if (CurrentFrame.mvpMapPoints[j]) cids[j] = NK;   // already claimed
        int32_t n = 0;
        gforb_throw(gforb_search_by_projection_keyframe(h_, &cur.set, &kf.set, 1, ids.data(), skip.data(), world.data(), mind.data(),
                                                        mpd.data(), CurrentFrame.mTcw, CurrentFrame.fx, CurrentFrame.fy, CurrentFrame.cx,
                                                        CurrentFrame.cy, th, ORBdist, mbCheckOrientation ? 1 : 0, cids.data(),
                                                        score.data(), &n, 0, nullptr),
                    "SearchByProjection(Current, KeyFrame)");
        for (int j = 0; j < CurrentFrame.N; j++) {
            if (cids[j] >= 0 && cids[j] < NK) CurrentFrame.mvpMapPoints[j] = pKF->mvpMapPoints[cids[j]];
            else if (cids[j] < 0) CurrentFrame.mvpMapPoints[j] = nullptr;
            CurrentFrame.mvpMatchScore[j] = score[j];
        }
        return n;
    }

    // ---- the overloads of the LocalMapping / LoopClosing threads
    // src/ORBmatcher.cc:1289-1424 (LoopClosing::ComputeSim3)
    int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12) {
        vpMatches12 = std::vector<MapPoint*>(pKF1->mvpMapPoints.size(), static_cast<MapPoint*>(nullptr));
        if (pKF1->N == 0 || pKF2->N == 0) return 0;
        ensure(std::max(pKF1->N, pKF2->N), pKF1->N, pKF2->mnScaleLevels, pKF2->mfScaleFactor);
        FlatFrame a(*pKF1), b(*pKF2);
        std::vector<int32_t> id1, id2, n1, n2, m(pKF1->N, -1);
        pair_inputs(*pKF1, id1, n1, true); pair_inputs(*pKF2, id2, n2, true);
        int32_t n = 0;
        gforb_throw(gforb_search_by_bow_keyframes(h_, &a.set, id1.data(), n1.data(), &b.set, id2.data(), n2.data(), 1, mfNNratio,
                                                  mbCheckOrientation ? 1 : 0, m.data(), &n, 0, nullptr), "SearchByBoW(KeyFrame, KeyFrame)");
        for (int i = 0; i < pKF1->N; i++) if (m[i] >= 0) vpMatches12[i] = pKF2->mvpMapPoints[m[i]];
        return n;
    }

    // src/ORBmatcher.cc:1426-1588 (LocalMapping::CreateNewMapPoints); F12: 3x3 CV_32F
    int SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, const gforb_cv::Mat& F12, std::vector<gforb_cv::KeyPoint>& vMatchedKeys1,
                               std::vector<gforb_cv::KeyPoint>& vMatchedKeys2, std::vector<std::pair<size_t, size_t>>& vMatchedPairs) {
        vMatchedKeys1.clear(); vMatchedKeys2.clear(); vMatchedPairs.clear();
        if (pKF1->N == 0 || pKF2->N == 0) return 0;
        ensure(std::max(pKF1->N, pKF2->N), pKF1->N, pKF2->mnScaleLevels, pKF2->mfScaleFactor);
        FlatFrame a(*pKF1), b(*pKF2);
        std::vector<int32_t> id1, id2, n1, n2, m(pKF1->N, -1);
        pair_inputs(*pKF1, id1, n1, false); pair_inputs(*pKF2, id2, n2, false);     // "there is already a MapPoint": any non-NULL
        float F[9];
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) F[r * 3 + c] = F12.at<float>(r, c);
        int32_t n = 0;
        gforb_throw(gforb_search_for_triangulation(h_, &a.set, id1.data(), n1.data(), &b.set, id2.data(), n2.data(), 1, F,
                                                   mbCheckOrientation ? 1 : 0, m.data(), &n, 0, nullptr), "SearchForTriangulation");
        for (int i = 0; i < pKF1->N; i++) {
            if (m[i] < 0) continue;
            vMatchedKeys1.push_back(pKF1->mvKeysUn[i]);
            vMatchedKeys2.push_back(pKF2->mvKeysUn[m[i]]);
            vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)m[i]));
        }
        return n;
    }

    // src/ORBmatcher.cc:1590-1708 (LocalMapping::SearchInNeighbors): the matching loop runs on the GPU for the whole list, the
    // map mutations are applied here in list order (a point an earlier entry made bad or added to pKF is skipped, :1614)
    int Fuse(KeyFrame* pKF, std::vector<MapPoint*>& vpMapPoints, float th = 2.5) {
        std::vector<int32_t> bi, bd;
        std::vector<uint8_t> skip(vpMapPoints.size(), 0);
        for (size_t i = 0; i < vpMapPoints.size(); i++) {
            MapPoint* p = vpMapPoints[i];
            skip[i] = (!p || p->isBad() || p->IsInKeyFrame(pKF)) ? 1 : 0;
        }
        if (!fuse_match(pKF, pKF->Tcw, 0, vpMapPoints, skip, th, bi, bd)) return 0;
        int nFused = 0;
        for (size_t i = 0; i < vpMapPoints.size(); i++) {
            MapPoint* pMP = vpMapPoints[i];
            if (!pMP) continue;
            if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
            if (bi[i] < 0 || bd[i] > TH_LOW) continue;
            MapPoint* pMPinKF = pKF->GetMapPoint(bi[i]);
            if (pMPinKF) {
                if (!pMPinKF->isBad()) pMP->Replace(pMPinKF);
            } else {
                pMP->AddObservation(pKF, bi[i]);
                pKF->AddMapPoint(pMP, bi[i]);
            }
            nFused++;
        }
        return nFused;
    }

    // src/ORBmatcher.cc:1710-1839 (LoopClosing::SearchAndFuse); Scw: 4x4 CV_32F
    int Fuse(KeyFrame* pKF, const gforb_cv::Mat& Scw, const std::vector<MapPoint*>& vpPoints, float th = 2.5) {
        float S[16];
        for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) S[r * 4 + c] = Scw.at<float>(r, c);
        const std::set<MapPoint*> spAlreadyFound = pKF->GetMapPoints();
        std::vector<int32_t> bi, bd;
        std::vector<uint8_t> skip(vpPoints.size(), 0);
        for (size_t i = 0; i < vpPoints.size(); i++) skip[i] = (vpPoints[i]->isBad() || spAlreadyFound.count(vpPoints[i])) ? 1 : 0;
        if (!fuse_match(pKF, S, 1, vpPoints, skip, th, bi, bd)) return 0;
        int nFused = 0;
        for (size_t i = 0; i < vpPoints.size(); i++) {
            MapPoint* pMP = vpPoints[i];
            if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
            if (bi[i] < 0 || bd[i] > TH_LOW) continue;
            MapPoint* pMPinKF = pKF->GetMapPoint(bi[i]);
            if (pMPinKF) {
                if (!pMPinKF->isBad()) pMPinKF->Replace(pMP);
            } else {
                pMP->AddObservation(pKF, bi[i]);
                pKF->AddMapPoint(pMP, bi[i]);
            }
            nFused++;
        }
        return nFused;
    }

    // src/ORBmatcher.cc:855-977 (LoopClosing::ComputeSim3 / DetectLoop); Scw: 4x4 CV_32F
    int SearchByProjection(KeyFrame* pKF, const gforb_cv::Mat& Scw, const std::vector<MapPoint*>& vpPoints,
                           std::vector<MapPoint*>& vpMatched, int th) {
        if (pKF->N == 0 || vpPoints.empty()) return 0;
        ensure(pKF->N, (int)vpPoints.size(), pKF->mnScaleLevels, pKF->mfScaleFactor);
        float S[16];
        for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) S[r * 4 + c] = Scw.at<float>(r, c);
        std::set<MapPoint*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
        spAlreadyFound.erase(static_cast<MapPoint*>(nullptr));
        const int L = (int)vpPoints.size();
        FlatFrame kf(*pKF);
        FlatPoints P(vpPoints);
        std::vector<int32_t> ids(L), matched(pKF->N, -1);
        for (int i = 0; i < L; i++) { ids[i] = i; P.skip[i] = (vpPoints[i]->isBad() || spAlreadyFound.count(vpPoints[i])) ? 1 : 0; }
        for (int j = 0; j < pKF->N; j++) if (vpMatched[j]) matched[j] = L;          // held by an earlier match
        int32_t n = 0;
        gforb_throw(gforb_search_by_projection_sim3(h_, &kf.set, 1, S, pKF->fx, pKF->fy, pKF->cx, pKF->cy, &P.set, ids.data(), (float)th,
                                                    matched.data(), &n, 0, nullptr), "SearchByProjection(KeyFrame, Scw)");
        for (int j = 0; j < pKF->N; j++) if (matched[j] >= 0 && matched[j] < L) vpMatched[j] = vpPoints[matched[j]];
        return n;
    }

    // src/ORBmatcher.cc:1841-2079 (LoopClosing::ComputeSim3); R12: 3x3, t12: 3x1 CV_32F
    int SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12, const gforb_cv::Mat& R12,
                     const gforb_cv::Mat& t12, float th) {
        if (pKF1->N == 0 || pKF2->N == 0) return 0;
        ensure(std::max(pKF1->N, pKF2->N), std::max(pKF1->N, pKF2->N), pKF2->mnScaleLevels, pKF2->mfScaleFactor);
        const int N1 = pKF1->N, N2 = pKF2->N;
        std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
        for (int i = 0; i < N1; i++) {
            MapPoint* pMP = vpMatches12[i];
            if (pMP) {
                vbAlreadyMatched1[i] = true;
                const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
                if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
            }
        }
        FlatFrame a(*pKF1), b(*pKF2);
        FlatPoints P1(pKF1->mvpMapPoints), P2(pKF2->mvpMapPoints);
        for (int i = 0; i < N1; i++) if (vbAlreadyMatched1[i]) P1.skip[i] = 1;
        for (int j = 0; j < N2; j++) if (vbAlreadyMatched2[j]) P2.skip[j] = 1;
        std::vector<int32_t> id2(N2, -1), m12(N1, -1);
        for (int j = 0; j < N2; j++) if (pKF2->mvpMapPoints[j]) id2[j] = j;
        float R[9], t[3];
        for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) R[r * 3 + c] = R12.at<float>(r, c); t[r] = t12.at<float>(r, 0); }
        int32_t n = 0;
        gforb_throw(gforb_search_by_sim3(h_, &a.set, &b.set, 1, &P1.set, &P2.set, id2.data(), pKF1->Tcw, pKF2->Tcw, &s12, R, t, pKF1->fx,
                                         pKF1->fy, pKF1->cx, pKF1->cy, th, m12.data(), &n, 0, nullptr), "SearchBySim3");
        for (int i = 0; i < N1; i++) if (m12[i] >= 0) vpMatches12[i] = pKF2->mvpMapPoints[m12[i]];
        return n;
    }

    float mfNNratio;
    bool mbCheckOrientation;

private:
    friend class Observability;
    // a list of MapPoint* -> GforbWorldPointSet (one problem); NULL and bad points are flagged in `skip`
    struct FlatPoints {
        std::vector<float> world, normal, mind, maxd;
        std::vector<uint8_t> desc, skip;
        int32_t count;
        GforbWorldPointSet set;
        explicit FlatPoints(const std::vector<MapPoint*>& v)
            : world(v.size() * 3, 0.f), normal(v.size() * 3, 0.f), mind(v.size(), 1.f), maxd(v.size(), 1.f), desc(v.size() * 32, 0),
              skip(v.size(), 0), count((int32_t)v.size()) {
            for (size_t i = 0; i < v.size(); i++) {
                const MapPoint* p = v[i];
                if (!p || p->isBad()) { skip[i] = 1; if (!p) continue; }
                std::memcpy(&world[i * 3], p->mWorldPos, 12); std::memcpy(&normal[i * 3], p->mNormalVector, 12);
                mind[i] = p->mfMinDistance; maxd[i] = p->mfMaxDistance;
                std::memcpy(&desc[i * 32], p->mDescriptor, 32);
            }
            set.world = world.data(); set.normal = normal.data(); set.min_distance = mind.data(); set.max_distance = maxd.data();
            set.descriptors = desc.data(); set.skip = skip.data(); set.counts = &count; set.mp_stride = (int32_t)std::max<size_t>(v.size(), 1);
        }
    };
    // ids (index of the keypoint when it holds a usable map point) and FeatureVector nodes of a keyframe pair member
    static void pair_inputs(const KeyFrame& KF, std::vector<int32_t>& ids, std::vector<int32_t>& nodes, bool dropBad) {
        ids.assign(KF.N, -1); nodes.assign(KF.N, -1);
        for (int i = 0; i < KF.N; i++) if (KF.mvpMapPoints[i] && !(dropBad && KF.mvpMapPoints[i]->isBad())) ids[i] = i;
        for (const auto& nd : KF.mFeatVec) for (unsigned i : nd.second) if ((int)i < KF.N) nodes[i] = (int32_t)nd.first;
    }
    template <class Vec>
    bool fuse_match(KeyFrame* pKF, const float* pose, int sim3, const Vec& pts, const std::vector<uint8_t>& skip, float th,
                    std::vector<int32_t>& bi, std::vector<int32_t>& bd) {
        const int L = (int)pts.size();
        bi.assign(L, -1); bd.assign(L, 0x7fffffff);
        if (pKF->N == 0 || L == 0) return false;
        ensure(pKF->N, L, pKF->mnScaleLevels, pKF->mfScaleFactor);
        FlatFrame kf(*pKF);
        FlatPoints P(std::vector<MapPoint*>(pts.begin(), pts.end()));
        for (int i = 0; i < L; i++) if (skip[i]) P.skip[i] = 1;
        int32_t n = 0;
        gforb_throw(gforb_fuse(h_, &kf.set, 1, pose, sim3, pKF->fx, pKF->fy, pKF->cx, pKF->cy, &P.set, th, bi.data(), bd.data(), &n, 0,
                               nullptr), "Fuse");
        return true;
    }                       // runActiveMapMatching drives the one-point matcher of this handle
    struct FlatFrame {          // Frame -> GforbFrameSet (one frame); keeps the staging alive for the call
        std::vector<uint8_t> desc;
        int32_t count;
        GforbFrameSet set;
        template <class FrameLike>
        explicit FlatFrame(const FrameLike& F) : desc((size_t)F.N * 32), count(F.N) {
            for (int i = 0; i < F.N; i++) std::memcpy(&desc[(size_t)i * 32], F.mDescriptors.ptr(i), 32);
            set.keypoints = reinterpret_cast<const GforbKeyPoint*>(F.mvKeysUn.data());
            set.descriptors = desc.data(); set.counts = &count; set.kp_stride = F.N;
            set.min_x = F.mnMinX; set.max_x = F.mnMaxX; set.min_y = F.mnMinY; set.max_y = F.mnMaxY;
        }
    };
    static std::vector<int32_t> ids_of(const Frame& F, const std::vector<MapPoint*>& pool) {
        // MapPoint* -> id: index into `pool` when the pointer is one of its entries, otherwise "some other map point"
        std::vector<int32_t> ids(F.N, -1);
        for (int i = 0; i < F.N; i++) {
            if (!F.mvpMapPoints[i]) continue;
            ids[i] = (int32_t)pool.size();
            for (size_t k = 0; k < pool.size(); k++) if (pool[k] == F.mvpMapPoints[i]) { ids[i] = (int32_t)k; break; }
        }
        return ids;
    }
    void ensure(int nKeys, int nQueries, int nlevels, float scaleFactor) {
        if (h_ && nKeys <= maxKeys_ && nQueries <= maxQ_ && nlevels == nlevels_ && scaleFactor == sf_) return;
        if (h_) { gforb_matcher_destroy(h_); h_ = nullptr; }
        maxKeys_ = std::max(maxKeys_, std::max(nKeys, 2048));
        maxQ_ = std::max(maxQ_, std::max(nQueries, 2048));
        nlevels_ = nlevels; sf_ = scaleFactor;
        GforbMatcherConfig c;
        std::memset(&c, 0, sizeof(c));
        c.max_batch = 1; c.max_keypoints = maxKeys_; c.max_queries = maxQ_; c.cand_pool = 0; c.nlevels = nlevels;
        c.scaleFactor = scaleFactor; c.device = device_;
        gforb_throw(gforb_matcher_create(&c, &h_), "ORBmatcher: create");
    }
    int device_;
    GforbMatcher* h_ = nullptr;
    int maxKeys_ = 0, maxQ_ = 0, nlevels_ = 0;
    float sf_ = 0;
};

// ------------------------------------------------------------------------------------------------ Observability
// Good-feature selection (DIMENSION_OF_STATE_MODEL = 7).  GoodPoint mirrors include/Util.hpp:64-92 (idx, obs_score,
// selected, upper_bound); the 7x7 block itself stays in HBM.
// include/Observability.h:94-102: namespace-scope constants, as the call sites name them (ORB_SLAM::FRAME_INFO_MATRIX)
enum { MAP_INFO_MATRIX = 0, FRAME_INFO_MATRIX = 1, MAP_HYBRID_MATRIX = 2, FRAME_HYBRID_MATRIX = 3, FRAME_STRIP_OBS_MATRIX = 4 };

struct GoodPoint {
    size_t idx = 0;
    double obs_score = -1, upper_bound = 0;
    bool selected = false;
};

class Observability {
public:
    // used by Tracking (src/Tracking.cc:146; k1 = k2 = 0)
    Observability(double fu, double fv, int nRows, int nCols, double Cx, double Cy, double k1, double k2, int device = 0) {
        init(fu, fv, nRows, nCols, Cx, Cy, k1, k2, 1.0, 1.0, device);
    }
    // used by the reference's tests / tools (fu = f / dx, fv = f / dy; include/Observability.h:117-160)
    Observability(double f, int nRows, int nCols, double Cx, double Cy, double k1, double k2, double dx, double dy,
                  int device = 0) {
        init(f / dx, f / dy, nRows, nCols, Cx, Cy, k1, k2, dx, dy, device);
    }
    ~Observability() { if (h_) gforb_obs_destroy(h_); }
    Observability(const Observability&) = delete;
    Observability& operator=(const Observability&) = delete;

    // include/Observability.h:222-259 / :261-295 (4x4 row-major float poses like the reference's cv::Mat)
    void updatePWLSVec(const double& time_prev, const gforb_cv::Mat& Tcw_prev, const double& time_cur,
                       const gforb_cv::Mat& Twc_cur) {
        float a[16], b[16];
        for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) { a[r * 4 + c] = Tcw_prev.at<float>(r, c); b[r * 4 + c] = Twc_cur.at<float>(r, c); }
        gforb_throw(gforb_obs_update_pwls(h_, time_prev, a, time_cur, b), "updatePWLSVec");
    }
    void predictPWLSVec(const double& dt, const size_t& num_seg_pred) {
        gforb_throw(gforb_obs_predict_pwls(h_, dt, (int)num_seg_pred), "predictPWLSVec");
    }
    void setState(const double* Xv13) { gforb_throw(gforb_obs_set_state(h_, Xv13), "setState"); }

    // the frame / map pool the next calls work on (the reference keeps pFrame / mMapPoints members)
    Frame* pFrame = nullptr;
    std::vector<MapPoint*>* mMapPoints = nullptr;
    std::vector<MapPoint*> mLeftMapPoints;            // include/Observability.h:731
    size_t mKineIdx = 0;
    long unsigned int mnFrameId = 0;                  // include/Observability.h:720 (set by Tracking before each call)

    // src/Observability.cc:646-804: per-match information blocks for the current pFrame (FRAME_INFO_MATRIX) or the
    // local map (MAP_INFO_MATRIX).  Time budgets are accepted and ignored (the device call is bounded by construction).
    bool runMatrixBuilding(const size_t mat_type, const double /*time_for_build*/, const bool /*with_multi_thread*/,
                           const bool check_viz = false) {
        double xv[13];
        if (gforb_obs_get_segment(h_, (int)mKineIdx, xv, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr) != GFORB_OK) return false;
        if (mat_type == FRAME_INFO_MATRIX) {
            if (!pFrame || pFrame->N == 0) return false;
            Frame& F = *pFrame;
            n_ = F.N;
            if (n_ > maxPoints_) return false;
            std::vector<float> world((size_t)n_ * 3, 0.f);
            std::vector<int32_t> oct(n_), sc(n_);
            std::vector<uint8_t> valid(n_, 0);
            for (int i = 0; i < n_; i++) {
                oct[i] = F.mvKeysUn[i].octave; sc[i] = F.mvpMatchScore[i];
                const MapPoint* p = F.mvpMapPoints[i];
                if (p && !F.mvbOutlier[i] && F.mvbCandidate[i]) { valid[i] = 1; std::memcpy(&world[(size_t)i * 3], p->mWorldPos, 12); }
            }
            blocks_.resize((size_t)n_ * 49); score_.resize(n_);
            int32_t cnt = n_;
            return gforb_obs_build_info_frame(h_, xv, world.data(), oct.data(), sc.data(), valid.data(), &cnt, n_, 1,
                                              F.mvLevelSigma2.data(), (int)F.mvLevelSigma2.size(), blocks_.data(),
                                              score_.data(), nullptr, 0, nullptr) == GFORB_OK;
        }
        if (mat_type == MAP_INFO_MATRIX) {
            if (!mMapPoints || mMapPoints->empty() || !pFrame) return false;
            n_ = (int)mMapPoints->size();
            if (n_ > maxPoints_) return false;
            std::vector<float> world((size_t)n_ * 3);
            std::vector<uint8_t> inView(n_);
            for (int i = 0; i < n_; i++) { std::memcpy(&world[(size_t)i * 3], (*mMapPoints)[i]->mWorldPos, 12); inView[i] = (*mMapPoints)[i]->mbTrackInView; }
            // points already refreshed at this frame are skipped (src/Observability.cc:586-587)
            std::vector<uint8_t> skip(n_);
            for (int i = 0; i < n_; i++) skip[i] = (*mMapPoints)[i]->updateAtFrameId == mnFrameId;
            blocks_.resize((size_t)n_ * 49); score_.resize(n_);
            std::vector<float> uv((size_t)n_ * 2);
            std::vector<double> hm((size_t)n_ * 14);
            int32_t cnt = n_;
            const bool ok = gforb_obs_build_info_map(h_, xv, world.data(), inView.data(), &cnt, n_, 1, check_viz ? 1 : 0,
                                                     pFrame->mnMinX, pFrame->mnMaxX, pFrame->mnMinY, pFrame->mnMaxY,
                                                     blocks_.data(), score_.data(), uv.data(), hm.data(), 0, nullptr) == GFORB_OK;
            for (int i = 0; ok && i < n_; i++) {
                MapPoint* p = (*mMapPoints)[i];
                if (skip[i] || score_[i] == -2.0) continue;                // refreshed earlier / not touched by the builder
                p->ObsScore = score_[i];
                if (score_[i] != 1.0) continue;
                p->u_proj = uv[(size_t)i * 2]; p->v_proj = uv[(size_t)i * 2 + 1];
                std::memcpy(p->H_meas, &hm[(size_t)i * 14], sizeof(p->H_meas));
                std::memcpy(p->ObsMat, &blocks_[(size_t)i * 49], sizeof(p->ObsMat));
                p->updateAtFrameId = mnFrameId;
            }
            return ok;
        }
        return false;
    }

    // src/Observability.cc:1249-1524.  mBaseInfoMat is a 7x7 row-major double[49] here (arma::mat in the reference); the
    // time budget is accepted and ignored; the hybrid matrix types need the 13-state model and return 0 matches.
    // The reference draws from the process-global rand(); `rand_seed` restarts glibc's generator for this call
    // (1 = a process that never called srand).
    int runActiveMapMatching(Frame* pF, const size_t mat_type, const double* mBaseInfoMat, const float th,
                             ORBmatcher& mORBMatcher, const int num_to_match, const double /*time_for_match*/ = 1e30,
                             uint32_t rand_seed = 1, double* curMatOut = nullptr) {
        if (pF == nullptr || mMapPoints == nullptr) return 0;
        if (mat_type != MAP_INFO_MATRIX && mat_type != FRAME_INFO_MATRIX) return 0;
        const int M = (int)mMapPoints->size();
        if (M == 0 || pF->N == 0) return 0;
        Frame& F = *pF;
        mORBMatcher.ensure(F.N, M, F.mnScaleLevels, F.mfScaleFactor);
        std::vector<float> u(M), v(M), vc(M);
        std::vector<int32_t> lvl(M);
        std::vector<uint8_t> fl(M), desc((size_t)M * 32);
        std::vector<double> osc(M), om((size_t)M * 49), hm((size_t)M * 14);
        for (int i = 0; i < M; i++) {
            const MapPoint* p = (*mMapPoints)[i];
            fl[i] = 0; osc[i] = -1.0;
            if (!p) continue;                                               // NULL entries are skipped (:1291)
            u[i] = p->mTrackProjX; v[i] = p->mTrackProjY; vc[i] = p->mTrackViewCos; lvl[i] = p->mnTrackScaleLevel;
            fl[i] = (uint8_t)((p->mbTrackInView ? 1 : 0) | (p->isBad() ? 2 : 0));
            std::memcpy(&desc[(size_t)i * 32], p->GetDescriptor(), 32);
            osc[i] = (p->updateAtFrameId == F.mnId) ? 1.0 : -1.0;           // :1297
            std::memcpy(&om[(size_t)i * 49], p->ObsMat, sizeof(p->ObsMat));
            std::memcpy(&hm[(size_t)i * 14], p->H_meas, sizeof(p->H_meas));
        }
        ORBmatcher::FlatFrame ff(F);
        std::vector<int32_t> ids = ORBmatcher::ids_of(F, *mMapPoints), score(F.mvpMatchScore.begin(), F.mvpMatchScore.end());
        std::vector<int32_t> left(M);
        int32_t mcount = M, nMatched = 0, nLeft = 0;
        double info[49];
        GforbMapPointSet mp = {u.data(), v.data(), lvl.data(), vc.data(), fl.data(), desc.data(), &mcount, M};
        gforb_throw(gforb_obs_active_match(h_, mORBMatcher.h_, &ff.set, &mp, 1, osc.data(), om.data(), hm.data(), mBaseInfoMat,
                                           F.mvLevelSigma2.data(), (int)F.mvLevelSigma2.size(), (int)mat_type, th,
                                           mORBMatcher.mfNNratio, num_to_match, &rand_seed, ids.data(), score.data(), &nMatched,
                                           left.data(), &nLeft, info, 0, nullptr), "runActiveMapMatching");
        for (int i = 0; i < F.N; i++) {
            if (ids[i] >= 0 && ids[i] < M) F.mvpMapPoints[i] = (*mMapPoints)[ids[i]];
            F.mvpMatchScore[i] = score[i];
        }
        for (int i = 0; i < nLeft; i++) mLeftMapPoints.push_back((*mMapPoints)[left[i]]);
        if (curMatOut) std::memcpy(curMatOut, info, sizeof(info));
        return nMatched;
    }

    // src/Observability.cc:808-1018: build the frame blocks, pool the valid matches in index order, pick k of them.
    // greedy_mtd: 1 baseline greedy, 2 lazier greedy, 3 automatic, 5 deletion.  rand() is an explicit stream here:
    // rng_seed = 1 reproduces a reference process that never called srand().
    bool setSelction_Number(const size_t num_good_inlier, const int greedy_mtd, const double time_for_select,
                            const double error_bound, Frame* pF, std::vector<GoodPoint>* mpVec) {
        if (!pF || !mpVec || num_good_inlier == 0) return false;          // :814-817
        pFrame = pF;
        const size_t N_frame = pF->mvpMapPoints.size();
        const double sampleScale = (double)(size_t)((float)N_frame / (float)num_good_inlier * std::log(1.0 / error_bound));  // :826
        if (!runMatrixBuilding(FRAME_INFO_MATRIX, time_for_select / 2, true, false)) return false;
        return select_pool(num_good_inlier, greedy_mtd, sampleScale, mpVec, pF);
    }
    // src/Observability.cc:1021-1247: the local-map variant.  mKineIdx = 1, runMatrixBuilding(MAP_INFO_MATRIX, check_viz),
    // pool = map points with ObsScore >= 0 in index order (GoodPoint::idx = index into *mapPoints), fixed
    // random_sample_scale = 6.0 (:1035).  greedy_mtd 3 with a pool of >= 2 * Greedy_Paral_TrigSz (1000) points that exceeds
    // Greedy_Paral_RelFac (1.2) * k by more than 10 is cut into threadNeeded = min(round(pool / 1000), mNumThreads) ranges
    // (:1085-1163): every range picks ceil(k / threadNeeded * 1.2) points by the automatic lazier greedy, the picks are
    // concatenated and the final k are picked from them.  In the reference those ranges run on std::threads that all draw
    // from the process-global rand() (a data race: its result depends on thread timing); here range i draws from its own
    // stream rng_seed + 1 + i and the final pass from rng_seed -- one GPU call for all ranges, one for the final pass.
    size_t mNumThreads = 8;                           // std::thread::hardware_concurrency() in the reference (include/Observability.h:138)
    bool setSelction_Number(const size_t num_good_inlier, const int greedy_mtd, const double time_for_select,
                            std::vector<MapPoint*>* mapPoints, std::vector<GoodPoint>* mpVec) {
        if (!mapPoints || !mpVec || time_for_select <= 0) return false;   // :1026-1027
        mKineIdx = 1;
        mMapPoints = mapPoints;
        const double sampleScale = 6.0;                                   // :1035
        if (!runMatrixBuilding(MAP_INFO_MATRIX, time_for_select / 2.0, true, true)) return false;
        std::vector<int> pool;
        for (int i = 0; i < n_; i++) if ((*mapPoints)[i] && (*mapPoints)[i]->ObsScore >= 0) pool.push_back(i);   // :1046-1058
        mpVec->clear();
        const size_t np = pool.size();
        if (np == 0) return true;
        size_t threadNeeded = 1;
        if (greedy_mtd == 3 && !((double)np - 1.2 * (float)num_good_inlier <= 10 || np < 2 * 1000)) {             // :1085-1092
            threadNeeded = (size_t)std::round((float)np / 1000.f);
            if (threadNeeded > mNumThreads) threadNeeded = mNumThreads;
        }
        if (threadNeeded <= 1) return select_indices(pool, num_good_inlier, greedy_mtd, sampleScale, rng_seed, mpVec);
        const size_t numSubsetParal = (size_t)std::ceil((float)num_good_inlier / (float)threadNeeded * 1.2);      // :1110
        const size_t numPoolParal = (size_t)std::ceil((float)np / (float)threadNeeded);                           // :1111
        // all ranges in ONE batched call: range i = pool entries [i * numPoolParal, min((i + 1) * numPoolParal, np))
        std::vector<double> pb(threadNeeded * numPoolParal * 49, 0.0), ps(threadNeeded * numPoolParal, -1.0);
        std::vector<int32_t> cnt(threadNeeded), idx(threadNeeded * numPoolParal), nsel(threadNeeded);
        std::vector<uint32_t> seeds(threadNeeded);
        for (size_t t = 0; t < np; t++) { std::memcpy(&pb[t * 49], &blocks_[(size_t)pool[t] * 49], 49 * sizeof(double)); ps[t] = score_[pool[t]]; }
        for (size_t i = 0; i < threadNeeded; i++) {
            const size_t st = i * numPoolParal, ed = std::min((i + 1) * numPoolParal, np);
            cnt[i] = (int32_t)(ed > st ? ed - st : 0);
            seeds[i] = rng_seed + 1 + (uint32_t)i;
        }
        if (gforb_obs_select(h_, pb.data(), ps.data(), cnt.data(), (int)numPoolParal, (int)threadNeeded, (int)numSubsetParal, 3,
                             sampleScale, seeds.data(), idx.data(), nullptr, nsel.data(), (int)numPoolParal, 0, nullptr) != GFORB_OK)
            return false;
        std::vector<int> merged;                                          // :1147-1153, parVec order
        for (size_t i = 0; i < threadNeeded; i++)
            for (int t = 0; t < nsel[i]; t++) merged.push_back(pool[i * numPoolParal + (size_t)idx[i * numPoolParal + t]]);
        return select_indices(merged, num_good_inlier, 3, sampleScale, rng_seed, mpVec);
    }

    // src/Observability.cc:2815-3029 on the pool built by the last runMatrixBuilding
    bool maxVolSelection_LazierGreedy(const size_t /*stIdx*/, const size_t /*edIdx*/, std::vector<GoodPoint>* subVec,
                                      const size_t mpLimit, const double sampleScale, const double /*max_time*/) {
        return select_pool(mpLimit, 2, sampleScale, subVec, nullptr);
    }
    uint32_t rng_seed = 1;

private:
    void init(double fu, double fv, int nRows, int nCols, double Cx, double Cy, double k1, double k2, double dx, double dy,
              int device) {
        if (gforb_device_count() <= 0) throw std::runtime_error("Observability: no CUDA device (the B200 path has no CPU fallback)");
        GforbObsConfig c;
        std::memset(&c, 0, sizeof(c));
        c.fu = fu; c.fv = fv; c.nRows = nRows; c.nCols = nCols; c.Cx = Cx; c.Cy = Cy; c.k1 = k1; c.k2 = k2; c.dx = dx; c.dy = dy;
        c.max_batch = 1; c.max_points = maxPoints_; c.device = device;
        gforb_throw(gforb_obs_create(&c, &h_), "Observability: create");
    }
    // one selection over the pool entries `pool` (indices into the arrays of the last runMatrixBuilding), appended to *out
    bool select_indices(const std::vector<int>& pool, size_t k, int mtd, double sampleScale, uint32_t seed, std::vector<GoodPoint>* out) {
        const int np = (int)pool.size();
        if (np == 0) return true;
        if (np > maxPoints_) return false;
        std::vector<double> pb((size_t)np * 49), ps(np), ld(std::max<size_t>(k, 1) + np);
        for (int t = 0; t < np; t++) { std::memcpy(&pb[(size_t)t * 49], &blocks_[(size_t)pool[t] * 49], 49 * sizeof(double)); ps[t] = score_[pool[t]]; }
        std::vector<int32_t> idx(np);
        int32_t cnt = np, nsel = 0;
        if (gforb_obs_select(h_, pb.data(), ps.data(), &cnt, np, 1, (int)k, mtd, sampleScale, &seed, idx.data(), ld.data(), &nsel, np, 0,
                             nullptr) != GFORB_OK) return false;
        for (int t = 0; t < nsel; t++) {
            GoodPoint g; g.idx = (size_t)pool[idx[t]]; g.obs_score = ps[idx[t]]; g.selected = true; g.upper_bound = ld[t];
            out->push_back(g);
        }
        return true;
    }
    bool select_pool(size_t k, int mtd, double sampleScale, std::vector<GoodPoint>* out, Frame* F) {
        if (n_ <= 0 || !out) return false;
        // lmkSelectPool = entries with a built block, in index order (:838-848)
        std::vector<int> pool;
        for (int i = 0; i < n_; i++) if (score_[i] >= 0) pool.push_back(i);
        const int np = (int)pool.size();
        out->clear();
        if (np == 0) return true;
        std::vector<double> pb((size_t)np * 49), ps(np), ld(std::max<size_t>(k, 1) + np);
        for (int t = 0; t < np; t++) { std::memcpy(&pb[(size_t)t * 49], &blocks_[(size_t)pool[t] * 49], 49 * sizeof(double)); ps[t] = score_[pool[t]]; }
        std::vector<int32_t> idx(np);
        int32_t cnt = np, nsel = 0;
        if (gforb_obs_select(h_, pb.data(), ps.data(), &cnt, np, 1, (int)k, mtd, sampleScale, &rng_seed, idx.data(), ld.data(),
                             &nsel, np, 0, nullptr) != GFORB_OK) return false;
        for (int t = 0; t < nsel; t++) {
            GoodPoint g; g.idx = (size_t)pool[idx[t]]; g.obs_score = ps[idx[t]]; g.selected = true; g.upper_bound = ld[t];
            out->push_back(g);
            if (F) F->mvbGoodFeature[g.idx] = true;                       // :1009-1010
        }
        return true;
    }
    GforbObs* h_ = nullptr;
    int maxPoints_ = 8192, n_ = 0;
    std::vector<double> blocks_, score_;
};

// ------------------------------------------------------------------------------------------------ Optimizer
// Optimizer::PoseOptimization(Frame*) (include/Optimizer.h, src/Optimizer.cc:279-413): motion-only bundle adjustment on the
// frame's matched map points; pFrame->mTcw and pFrame->mvbOutlier are updated, returns nInitialCorrespondences - nBad.
class Optimizer {
public:
    static int PoseOptimization(Frame* pFrame, int device = 0) {
        if (!pFrame) return 0;
        Frame& F = *pFrame;
        const int N = (int)F.mvpMapPoints.size();
        if (N == 0) return 0;
        GforbPoseOpt* h = nullptr;
        gforb_throw(gforb_poseopt_create(1, N, device, &h), "PoseOptimization: create");
        std::vector<GforbKeyPoint> kps(N);
        std::vector<int32_t> mp(N);
        std::vector<float> world((size_t)N * 3, 0.f), inv(F.mvLevelSigma2.size());
        std::vector<uint8_t> out(N);
        for (int i = 0; i < N; i++) {
            std::memcpy(&kps[i], &F.mvKeysUn[i], sizeof(GforbKeyPoint));
            mp[i] = F.mvpMapPoints[i] ? i : -1;
            if (F.mvpMapPoints[i]) std::memcpy(&world[(size_t)i * 3], F.mvpMapPoints[i]->mWorldPos, 12);
            out[i] = F.mvbOutlier[i] ? 1 : 0;
        }
        for (size_t l = 0; l < inv.size(); l++) inv[l] = 1.0f / F.mvLevelSigma2[l];            // Frame::mvInvLevelSigma2 (src/Frame.cc:109-112)
        int32_t cnt = N, ninl = 0;
        const int rc = gforb_pose_optimization(h, F.mTcw, kps.data(), mp.data(), world.data(), &cnt, N, 1, inv.data(), (int)inv.size(),
                                               F.fx, F.fy, F.cx, F.cy, out.data(), &ninl, 0, nullptr);
        gforb_poseopt_destroy(h);
        gforb_throw(rc, "PoseOptimization");
        for (int i = 0; i < N; i++) if (F.mvpMapPoints[i]) F.mvbOutlier[i] = out[i] != 0;
        return ninl;
    }
};

}  // namespace ORB_SLAM
#endif  // GFORB_ORB_SLAM_HPP
