// GPU parity test of the C++ drop-in shim (include/gforb/orb_slam.hpp over libgforb.so) against the CPU oracle
// (oracle/oracle.hpp, liboracle.so): it reads like a test of the reference's own classes.
//   ORBextractor::operator()  ->  keypoints + descriptors bit-exact
//   ORBmatcher::SearchForInitialization / SearchByProjection / DescriptorDistance -> matches bit-exact
//   Observability::setSelction_Number (lazier greedy) -> selected subset bit-exact, log-det within 1e-5
// Built by __graft_entry__.build(); run by tests/test_gpu_shim.py (needs a CUDA device).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <set>
#include <vector>

#include "gforb/orb_slam.hpp"
#include "../../oracle/oracle.hpp"

static int g_fail = 0;
#define CHECK(cond, ...)                                            \
    do {                                                            \
        if (!(cond)) { g_fail++; printf("FAIL %s:%d: ", __FILE__, __LINE__); printf(__VA_ARGS__); printf("\n"); } \
    } while (0)

static uint64_t sm_state;
static uint64_t splitmix() {
    uint64_t z = (sm_state += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// blocky texture + noise, shifted by (dx, dy): corner-rich, and consecutive frames overlap
static void make_frame(std::vector<uint8_t>& img, int w, int h, int dx, int dy, uint64_t noiseSeed) {
    const int MW = w + 64, MH = h + 64;
    static std::vector<uint8_t> master;
    if (master.empty()) {
        sm_state = 12345;
        std::vector<uint8_t> base((MW / 4 + 2) * (MH / 4 + 2));
        for (auto& b : base) b = (uint8_t)(splitmix() >> 56);
        std::vector<int> up(MW * MH);
        for (int y = 0; y < MH; y++) for (int x = 0; x < MW; x++) up[y * MW + x] = base[(y / 4) * (MW / 4 + 2) + x / 4];
        master.resize(MW * MH);
        for (int y = 0; y < MH; y++)
            for (int x = 0; x < MW; x++) {
                int s = 0, n = 0;
                for (int v = -1; v <= 1; v++) for (int u = -1; u <= 1; u++) {
                    int yy = y + v, xx = x + u;
                    if (yy >= 0 && yy < MH && xx >= 0 && xx < MW) { s += up[yy * MW + xx]; n++; }
                }
                master[y * MW + x] = (uint8_t)(s / n);
            }
    }
    sm_state = noiseSeed;
    img.resize((size_t)w * h);
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            int v = master[(y + dy) * MW + x + dx] + (int)((splitmix() >> 56) % 7) - 3;
            img[(size_t)y * w + x] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
}

using namespace ORB_SLAM;

int main() {
    const int W = 640, H = 480;
    std::vector<uint8_t> im1, im2;
    make_frame(im1, W, H, 10, 8, 1);
    make_frame(im2, W, H, 13, 9, 2);

    // ---- extraction
    ORBextractor ex(1000, 1.2f, 8, ORBextractor::FAST_SCORE, 20);
    oracle::ORBextractor oex(1000, 1.2f, 8, oracle::FAST_SCORE, 20);
    std::vector<gforb_cv::KeyPoint> k1, k2;
    gforb_cv::Mat d1, d2, mask;
    ex(gforb_cv::Mat(H, W, im1.data(), W), mask, k1, d1);
    ex(gforb_cv::Mat(H, W, im2.data(), W), mask, k2, d2);
    std::vector<oracle::KeyPoint> ok1, ok2;
    std::vector<uint8_t> od1, od2;
    CHECK(oex(im1.data(), W, H, W, ok1, od1) == 0, "oracle extract 1");
    CHECK(oex(im2.data(), W, H, W, ok2, od2) == 0, "oracle extract 2");
    CHECK(k1.size() == ok1.size() && k2.size() == ok2.size(), "keypoint counts %zu/%zu %zu/%zu", k1.size(), ok1.size(), k2.size(), ok2.size());
    CHECK(k1.size() > 500, "too few keypoints: %zu", k1.size());
    if (k1.size() == ok1.size()) {
        CHECK(std::memcmp(k1.data(), ok1.data(), k1.size() * 28) == 0, "keypoints differ");
        bool same = true;
        for (size_t i = 0; i < k1.size(); i++) same = same && std::memcmp(d1.ptr((int)i), &od1[i * 32], 32) == 0;
        CHECK(same, "descriptors differ");
    }
    { gforb_cv::Mat e; std::vector<gforb_cv::KeyPoint> ke(3); ex(gforb_cv::Mat(), mask, ke, e); CHECK(ke.size() == 3, "empty image must return silently"); }

    // ---- matching
    ORBmatcher m(0.9f, true);
    oracle::ORBmatcher om(0.9f, true);
    Frame F1(k1, d1, W, H), F2(k2, d2, W, H);
    oracle::Frame oF1, oF2;
    oF1.init(ok1.data(), od1.data(), (int)ok1.size(), W, H, 8, 1.2f);
    oF2.init(ok2.data(), od2.data(), (int)ok2.size(), W, H, 8, 1.2f);
    std::vector<gforb_cv::Point2f> prev(F1.N);
    std::vector<float> oprev((size_t)F1.N * 2);
    for (int i = 0; i < F1.N; i++) { prev[i] = k1[i].pt; oprev[2 * i] = k1[i].pt.x; oprev[2 * i + 1] = k1[i].pt.y; }
    std::vector<int> m12, om12;
    const int nm = m.SearchForInitialization(F1, F2, prev, m12, 100);
    const int onm = om.SearchForInitialization(oF1, oF2, oprev, om12, 100);
    CHECK(nm == onm && m12 == om12, "SearchForInitialization: %d vs %d matches", nm, onm);
    CHECK(nm > 20, "too few matches: %d", nm);
    bool prevSame = true;
    for (int i = 0; i < F1.N; i++) prevSame = prevSame && prev[i].x == oprev[2 * i] && prev[i].y == oprev[2 * i + 1];
    CHECK(prevSame, "vbPrevMatched differs");
    CHECK(ORBmatcher::DescriptorDistance(d1.row(0), d2.row(0)) == oracle::ORBmatcher::DescriptorDistance(&od1[0], &od2[0]), "DescriptorDistance");

    // SearchByProjection(F, map points): project frame-1 keypoints (jittered) into frame 2
    std::vector<MapPoint> mps(F1.N);
    std::vector<MapPoint*> vp(F1.N);
    std::vector<oracle::MapPointView> omp(F1.N);
    sm_state = 77;
    for (int i = 0; i < F1.N; i++) {
        MapPoint& p = mps[i];
        p.mTrackProjX = k1[i].pt.x - 3.0f + (float)(splitmix() % 5) - 2.0f;
        p.mTrackProjY = k1[i].pt.y - 1.0f + (float)(splitmix() % 5) - 2.0f;
        p.mnTrackScaleLevel = k1[i].octave;
        p.mTrackViewCos = (splitmix() & 1) ? 0.999f : 0.9f;
        p.mbTrackInView = (splitmix() % 10) != 0;
        p.mbBad = (splitmix() % 20) == 0;
        std::memcpy(p.mDescriptor, d1.ptr(i), 32);
        vp[i] = &p;
        oracle::MapPointView& o = omp[i];
        o.mTrackProjX = p.mTrackProjX; o.mTrackProjY = p.mTrackProjY; o.mnTrackScaleLevel = p.mnTrackScaleLevel;
        o.mTrackViewCos = p.mTrackViewCos; o.mbTrackInView = p.mbTrackInView; o.isBad = p.mbBad;
        std::memcpy(o.desc, p.mDescriptor, 32);
    }
    {   // SearchByProjection_GoodFeature (src/ORBmatcher.cc:51-140) and _Budget (:276-379) on copies of the frame
        Frame Fg = F2;
        oracle::Frame oFg = oF2;
        Fg.mnId = 5;
        std::vector<uint8_t> good(F1.N);
        for (int i = 0; i < F1.N; i++) { good[i] = (splitmix() % 3) == 0; mps[i].goodAtFrameId = good[i] ? 5 : 4; }
        const int ng = m.SearchByProjection_GoodFeature(Fg, vp, 1.0f, 3.0f);
        const int ong = om.SearchByProjection_GoodFeature(oFg, omp, good.data(), 1.0f, 3.0f);
        CHECK(ng == ong && ng > 50, "SearchByProjection_GoodFeature: %d vs %d", ng, ong);
        bool ag = true;
        for (int i = 0; i < Fg.N; i++) {
            const int id = Fg.mvpMapPoints[i] ? (int)(Fg.mvpMapPoints[i] - &mps[0]) : -1;
            ag = ag && id == oFg.mvpMapPoints[i] && Fg.mvpMatchScore[i] == F2.mvpMatchScore[i];
        }
        CHECK(ag, "SearchByProjection_GoodFeature associations differ");
        Frame Fb = F2;
        CHECK(m.SearchByProjection_Budget(Fb, vp, 3.0f, 0.0) == 0, "SearchByProjection_Budget with a spent budget");
    }
    const int np = m.SearchByProjection(F2, vp, 3.0f);
    const int onp = om.SearchByProjection(oF2, omp, 3.0f);
    CHECK(np == onp, "SearchByProjection: %d vs %d", np, onp);
    bool assoc = true;
    for (int i = 0; i < F2.N; i++) {
        const int id = F2.mvpMapPoints[i] ? (int)(F2.mvpMapPoints[i] - &mps[0]) : -1;
        assoc = assoc && id == oF2.mvpMapPoints[i] && F2.mvpMatchScore[i] == oF2.mvpMatchScore[i];
    }
    CHECK(assoc, "SearchByProjection associations differ");
    CHECK(np > 50, "too few projection matches: %d", np);

    // ---- good-feature selection on the matched frame (tools/simu_greedy.cc camera)
    const double f = 5.1369248, dxy = 0.01123325985;
    Observability ob(f, 480, 752, 367.215, 248.375, 0.0, 0.0, dxy, dxy);
    oracle::Observability oob(f, 480, 752, 367.215, 248.375, 0.0, 0.0, dxy, dxy);
    const double Xv[13] = {0.3, -0.1, 1.0, 0.9992, 0.0131, 0.0314, -0.0209, 3.0, -1.0, 10.0, 0.2618, 0.6283, -0.4189};
    ob.setState(Xv); ob.predictPWLSVec(0.1, 1);
    std::memcpy(oob.Xv, Xv, sizeof(Xv)); oob.predictPWLSVec(0.1, 1);
    sm_state = 99;
    const int n2 = F2.N;
    std::vector<float> world((size_t)n2 * 3);
    std::vector<int> oct(n2), sc(n2);
    std::vector<uint8_t> valid(n2, 0);
    for (int i = 0; i < n2; i++) {
        MapPoint* p = F2.mvpMapPoints[i];
        if (!p) continue;
        p->mWorldPos[0] = (float)((splitmix() >> 11) * (1.0 / 9007199254740992.0) * 4.0 - 2.0);
        p->mWorldPos[1] = (float)((splitmix() >> 11) * (1.0 / 9007199254740992.0) * 4.0 - 2.0);
        p->mWorldPos[2] = (float)((splitmix() >> 11) * (1.0 / 9007199254740992.0) * 6.0 + 4.0);
        std::memcpy(&world[(size_t)i * 3], p->mWorldPos, 12);
        valid[i] = 1;
    }
    for (int i = 0; i < n2; i++) { oct[i] = k2[i].octave; sc[i] = F2.mvpMatchScore[i]; }
    std::vector<GoodPoint> good;
    const size_t K = 40;
    CHECK(ob.setSelction_Number(K, 2, 1.0, 0.1, &F2, &good), "setSelction_Number failed");
    // oracle: same pool
    std::vector<double> blocks((size_t)n2 * 49), oscore(n2);
    oob.batchInfoMat_Frame(world.data(), oct.data(), sc.data(), valid.data(), n2, F2.mvLevelSigma2.data(), blocks.data(), oscore.data(), nullptr);
    std::vector<int> pool;
    for (int i = 0; i < n2; i++)
        if (oscore[i] >= 0) {
            oracle::GoodPoint g; g.idx = i; g.obs_score = oscore[i]; g.selected = false; g.upper_bound = 0;
            std::memcpy(g.obs_block, &blocks[(size_t)i * 49], sizeof(g.obs_block));
            oob.lmkSelectPool.push_back(g); pool.push_back(i);
        }
    const double sampleScale = (double)(size_t)((float)n2 / (float)K * std::log(1.0 / 0.1));
    std::vector<oracle::GoodPoint> ogood;
    srand(1);
    oob.maxVolSelection_LazierGreedy(0, oob.lmkSelectPool.size(), &ogood, K, sampleScale, 1e9);
    CHECK(good.size() == ogood.size() && good.size() == K, "selected %zu vs %zu (pool %zu)", good.size(), ogood.size(), pool.size());
    bool subset = good.size() == ogood.size(), ldok = subset;
    for (size_t t = 0; subset && t < good.size(); t++) {
        subset = good[t].idx == ogood[t].idx;
        ldok = ldok && std::fabs(good[t].upper_bound - ogood[t].upper_bound) <= 1e-5 * std::fabs(ogood[t].upper_bound);
    }
    CHECK(subset, "good-feature subset differs");
    CHECK(ldok, "log-det trajectory differs by more than 1e-5 relative");

    // ---- active map matching (the default GF mode): fresh frame 2, every map point refreshed by the map builder
    int nActive = 0;
    {
        Frame F3(k2, d2, W, H);
        F3.mnId = 42;
        oracle::Frame oF3;
        oF3.init(ok2.data(), od2.data(), (int)ok2.size(), W, H, 8, 1.2f);
        sm_state = 1234;
        const int M = (int)mps.size();
        std::vector<float> mw((size_t)M * 3);
        std::vector<uint8_t> tv(M);
        for (int i = 0; i < M; i++) {
            MapPoint& p = mps[i];
            p.mWorldPos[0] = (float)((splitmix() >> 11) * (1.0 / 9007199254740992.0) * 4.0 - 2.0);
            p.mWorldPos[1] = (float)((splitmix() >> 11) * (1.0 / 9007199254740992.0) * 4.0 - 2.0);
            p.mWorldPos[2] = (float)((splitmix() >> 11) * (1.0 / 9007199254740992.0) * 6.0 + 4.0);
            p.updateAtFrameId = 7;
            std::memcpy(&mw[(size_t)i * 3], p.mWorldPos, 12);
            tv[i] = p.mbTrackInView;
        }
        ob.pFrame = &F3; ob.mMapPoints = &vp; ob.mnFrameId = F3.mnId;
        CHECK(ob.runMatrixBuilding(ORB_SLAM::MAP_INFO_MATRIX, 1.0, true, false), "runMatrixBuilding(MAP) failed");
        std::vector<double> mb((size_t)M * 49), ms(M), mh((size_t)M * 14);
        oob.batchInfoMat_Map(mw.data(), tv.data(), M, false, 0, W, 0, H, mb.data(), ms.data(), nullptr, mh.data());
        std::vector<uint8_t> inPool(M);
        bool built = true;
        for (int i = 0; i < M; i++) {
            inPool[i] = ms[i] == 1.0;
            built = built && ((mps[i].updateAtFrameId == F3.mnId) == (ms[i] == 1.0));
            if (ms[i] == 1.0) built = built && std::memcmp(mps[i].ObsMat, &mb[(size_t)i * 49], 392) == 0 &&
                                      std::memcmp(mps[i].H_meas, &mh[(size_t)i * 14], 112) == 0;
        }
        CHECK(built, "map information blocks differ");
        double base[49], ocur[49], cur[49];
        for (int i = 0; i < 49; i++) base[i] = (i % 8 == 0) ? 1e-5 : 0.0;
        std::memcpy(ocur, base, sizeof(base));
        ORBmatcher m3(0.8f, true);
        oracle::ORBmatcher om3(0.8f, true);
        nActive = ob.runActiveMapMatching(&F3, ORB_SLAM::MAP_INFO_MATRIX, base, 3.0f, m3, 60, 1e30, 1, cur);
        std::vector<int> oleft;
        srand(1);
        const int onActive = oracle::runActiveMapMatching(oF3, omp, inPool.data(), mb.data(), mh.data(), ocur, 3.0f, om3, 60, oleft);
        CHECK(nActive == onActive, "runActiveMapMatching: %d vs %d", nActive, onActive);
        CHECK(nActive > 20, "too few active matches: %d", nActive);
        bool a3 = true;
        for (int i = 0; i < F3.N; i++) {
            const int id = F3.mvpMapPoints[i] ? (int)(F3.mvpMapPoints[i] - &mps[0]) : -1;
            a3 = a3 && id == oF3.mvpMapPoints[i] && F3.mvpMatchScore[i] == oF3.mvpMatchScore[i];
        }
        CHECK(a3, "active matching associations differ");
        bool l3 = ob.mLeftMapPoints.size() == oleft.size();
        for (size_t i = 0; l3 && i < oleft.size(); i++) l3 = (int)(ob.mLeftMapPoints[i] - &mps[0]) == oleft[i];
        CHECK(l3, "mLeftMapPoints differs (%zu vs %zu)", ob.mLeftMapPoints.size(), oleft.size());
        bool c3 = true;
        for (int i = 0; i < 49; i++) c3 = c3 && std::fabs(cur[i] - ocur[i]) <= 1e-12 * std::fabs(ocur[i]);
        CHECK(c3, "accumulated information matrix differs");
    }

    // ---- Frame::UndistortKeyPoints through the shim vs the oracle's cv::undistortPoints restatement (EuRoC cam0 model)
    {
        const std::vector<float> dist = {-0.28340811f, 0.07395907f, 0.00019359f, 1.76187114e-05f};
        Frame Fu = F1;
        Fu.fx = 458.654f; Fu.fy = 457.296f; Fu.cx = 367.215f; Fu.cy = 248.375f;
        std::vector<gforb_cv::KeyPoint> un;
        m.UndistortKeyPoints(Fu, k1, dist, un);
        std::vector<float> src(2 * k1.size()), dst(2 * k1.size());
        for (size_t i = 0; i < k1.size(); i++) { src[2 * i] = k1[i].pt.x; src[2 * i + 1] = k1[i].pt.y; }
        oracle::undistortPoints(src.data(), (int)k1.size(), Fu.fx, Fu.fy, Fu.cx, Fu.cy, dist.data(), 4, dst.data());
        bool u1 = un.size() == k1.size(), moved = false;
        for (size_t i = 0; u1 && i < k1.size(); i++) {
            u1 = un[i].pt.x == dst[2 * i] && un[i].pt.y == dst[2 * i + 1] && un[i].octave == k1[i].octave && un[i].angle == k1[i].angle &&
                 un[i].response == k1[i].response && un[i].size == k1[i].size;
            moved = moved || std::fabs(un[i].pt.x - k1[i].pt.x) > 1.0f;
        }
        CHECK(u1 && moved, "UndistortKeyPoints differs from the oracle");
    }

    // ---- MapPoint::ComputeDistinctiveDescriptors through the shim vs the oracle (observations = jittered copies)
    {
        const int nmp = 40;
        std::vector<MapPoint> dmp(nmp);
        std::vector<MapPoint*> dvp(nmp);
        std::vector<std::vector<uint8_t>> obsd(nmp);
        sm_state = 4242;
        for (int i = 0; i < nmp; i++) {
            dvp[i] = &dmp[i];
            std::memset(dmp[i].mDescriptor, 0x5a, 32);
            const int nobs = i % 9;                                   // 0 .. 8 observations
            obsd[i].resize((size_t)nobs * 32);
            for (int r = 0; r < nobs; r++) {
                std::memcpy(&obsd[i][(size_t)r * 32], d1.ptr(i), 32);
                const int flips = (int)(splitmix() % 40);
                for (int f = 0; f < flips; f++) { const int b = (int)(splitmix() % 256); obsd[i][(size_t)r * 32 + (b >> 3)] ^= (uint8_t)(1u << (b & 7)); }
            }
        }
        m.ComputeDistinctiveDescriptors(dvp, obsd);
        bool dd = true;
        for (int i = 0; i < nmp; i++) {
            const int nobs = i % 9;
            const int ob = oracle::distinctiveDescriptor(obsd[i].data(), nobs);
            if (ob < 0) { for (int q = 0; q < 32; q++) dd = dd && dmp[i].mDescriptor[q] == 0x5a; }
            else dd = dd && std::memcmp(dmp[i].mDescriptor, &obsd[i][(size_t)ob * 32], 32) == 0;
        }
        CHECK(dd, "ComputeDistinctiveDescriptors differs from the oracle");
    }

    // ---- ORBmatcher::SearchByProjection(F1, F2, windowSize, out) (TrackPreviousFrame) and Optimizer::PoseOptimization through the
    //      shim vs the oracle: frame-1 keypoints become map points that project near their frame-2 positions
    int nPrev = 0, nInl = 0;
    {
        const float fx = 458.0f, fy = 457.0f, cx = 376.0f, cy = 240.0f;
        std::vector<MapPoint> wp(F1.N);
        std::vector<float> world((size_t)F1.N * 3);
        sm_state = 909;
        for (int i = 0; i < F1.N; i++) {
            const float z = 2.0f + (float)(splitmix() % 600) * 0.01f;
            const float u = k1[i].pt.x - 3.0f + (float)(splitmix() % 3) - 1.0f, v = k1[i].pt.y - 1.0f + (float)(splitmix() % 3) - 1.0f;
            wp[i].mWorldPos[0] = (u - cx) / fx * z; wp[i].mWorldPos[1] = (v - cy) / fy * z; wp[i].mWorldPos[2] = z;
            wp[i].mbBad = (splitmix() % 25) == 0;
            std::memcpy(&world[(size_t)i * 3], wp[i].mWorldPos, 12);
        }
        Frame Fa = F1, Fb = F2;
        oracle::Frame oFa = oF1, oFb = oF2;
        for (int i = 0; i < Fa.N; i++) {
            const bool has = (splitmix() % 7) != 0;
            Fa.mvpMapPoints[i] = has ? &wp[i] : nullptr;
            oFa.mvpMapPoints[i] = (has && !wp[i].mbBad) ? i : -1;
        }
        for (int j = 0; j < Fb.N; j++) {                       // F2 arrives holding a few points already (some shared with F1)
            Fb.mvpMapPoints[j] = nullptr; oFb.mvpMapPoints[j] = -1; Fb.mvpMatchScore[j] = 999; oFb.mvpMatchScore[j] = 999;
            if ((splitmix() % 12) == 0) { const int i = (int)(splitmix() % (uint64_t)Fa.N); Fb.mvpMapPoints[j] = &wp[i]; oFb.mvpMapPoints[j] = i; }
        }
        // the product maps pointer identity to ids itself; the oracle ids above use the F1 index of the shared point only when F1
        // holds it -- align the oracle with the shim's convention (a point only F2 holds: F1.N + j)
        for (int j = 0; j < Fb.N; j++)
            if (Fb.mvpMapPoints[j]) {
                const int i = (int)(Fb.mvpMapPoints[j] - &wp[0]);
                oFb.mvpMapPoints[j] = (Fa.mvpMapPoints[i] == Fb.mvpMapPoints[j]) ? i : Fa.N + j;
            }
        Fb.fx = fx; Fb.fy = fy; Fb.cx = cx; Fb.cy = cy; oFb.fx = fx; oFb.fy = fy; oFb.cx = cx; oFb.cy = cy;
        ORBmatcher mp(0.9f, true);
        oracle::ORBmatcher omp2(0.9f, true);
        std::vector<MapPoint*> out2;
        std::vector<int> oout2;
        nPrev = mp.SearchByProjection(Fa, Fb, 15, out2);
        const int onPrev = omp2.SearchByProjection(oFa, oFb, world.data(), 15, oout2);
        CHECK(nPrev == onPrev && nPrev > 50, "SearchByProjection(F1, F2, window): %d vs %d", nPrev, onPrev);
        bool pa = out2.size() == oout2.size();
        for (size_t j = 0; pa && j < out2.size(); j++) {
            int id = -1;
            if (out2[j]) { const int i = (int)(out2[j] - &wp[0]); id = (Fa.mvpMapPoints[i] == out2[j]) ? i : Fa.N + (int)j; }
            pa = id == oout2[j] && Fb.mvpMatchScore[j] == oFb.mvpMatchScore[j];
        }
        CHECK(pa, "SearchByProjection(F1, F2, window) associations / scores differ");
        // PoseOptimization on the matched frame from a perturbed pose
        Frame Fp = Fb;
        Fp.mvpMapPoints = out2;
        Fp.mTcw[3] = 0.02f; Fp.mTcw[7] = -0.01f; Fp.mTcw[11] = 0.03f;
        std::vector<float> xy((size_t)Fp.N * 2), wpos((size_t)Fp.N * 3, 0.f), inv(Fp.mvLevelSigma2.size());
        std::vector<int> oct(Fp.N), ids(Fp.N);
        std::vector<uint8_t> oflag(Fp.N, 0);
        for (int j = 0; j < Fp.N; j++) {
            xy[2 * j] = Fp.mvKeysUn[j].pt.x; xy[2 * j + 1] = Fp.mvKeysUn[j].pt.y; oct[j] = Fp.mvKeysUn[j].octave;
            ids[j] = out2[j] ? j : -1;
            if (out2[j]) std::memcpy(&wpos[(size_t)j * 3], out2[j]->mWorldPos, 12);
        }
        for (size_t l = 0; l < inv.size(); l++) inv[l] = 1.0f / Fp.mvLevelSigma2[l];
        float oT[16];
        std::memcpy(oT, Fp.mTcw, sizeof(oT));
        const int onInl = oracle::PoseOptimization(oT, Fp.N, xy.data(), oct.data(), ids.data(), wpos.data(), inv.data(), fx, fy, cx, cy, oflag.data());
        nInl = Optimizer::PoseOptimization(&Fp);
        CHECK(nInl == onInl && nInl > 30, "PoseOptimization inliers: %d vs %d", nInl, onInl);
        bool fl = true, ps = true;
        for (int j = 0; j < Fp.N; j++) if (out2[j]) fl = fl && (Fp.mvbOutlier[j] ? 1 : 0) == oflag[j];
        for (int e = 0; e < 16; e++) ps = ps && std::fabs(Fp.mTcw[e] - oT[e]) <= 1e-5f * std::fabs(oT[e]) + 1e-6f;
        CHECK(fl, "PoseOptimization outlier flags differ");
        CHECK(ps, "PoseOptimization pose differs by more than 1e-5 relative");
    }

    // ---- the Tracking-side keyframe overloads through the shim vs the oracle (TrackReferenceKeyFrame / Relocalization):
    //      ORBmatcher::SearchByBoW(KeyFrame*, Frame&, out) and SearchByProjection(CurrentFrame, KeyFrame*, sAlreadyFound, th, ORBdist)
    int nBow = 0, nKf = 0;
    {
        const float fx = 458.0f, fy = 457.0f, cx = 376.0f, cy = 240.0f;
        std::vector<MapPoint> wp(F1.N);
        std::vector<float> world((size_t)F1.N * 3), mind(F1.N);
        std::vector<uint8_t> mpd((size_t)F1.N * 32), skip(F1.N, 0);
        std::set<MapPoint*> found;
        sm_state = 4242;
        Frame Fk = F1, Fc = F2;
        oracle::Frame oK = oF1, oC = oF2;
        float sc[8]; sc[0] = 1.0f; for (int l = 1; l < 8; l++) sc[l] = sc[l - 1] * 1.2f;
        for (int i = 0; i < F1.N; i++) {
            const float z = 2.0f + (float)(splitmix() % 600) * 0.01f;
            const float u = k1[i].pt.x - 3.0f, v = k1[i].pt.y - 1.0f;        // frame 2 is frame 1 shifted by (3, 1)
            wp[i].mWorldPos[0] = (u - cx) / fx * z; wp[i].mWorldPos[1] = (v - cy) / fy * z; wp[i].mWorldPos[2] = z;
            const float d3 = std::sqrt(wp[i].mWorldPos[0] * wp[i].mWorldPos[0] + wp[i].mWorldPos[1] * wp[i].mWorldPos[1] + z * z);
            wp[i].mfMinDistance = d3 / (sc[k1[i].octave] * (0.95f + 0.01f * (float)(splitmix() % 10)));
            std::memcpy(wp[i].mDescriptor, F1.mDescriptors.ptr(i), 32);
            if ((splitmix() % 5) == 0) wp[i].mDescriptor[splitmix() % 32] ^= 0x22;
            wp[i].mbBad = (splitmix() % 30) == 0;
            const bool has = (splitmix() % 8) != 0;
            Fk.mvpMapPoints[i] = has ? &wp[i] : nullptr;
            if (has && (splitmix() % 9) == 0) found.insert(&wp[i]);
            oK.mvpMapPoints[i] = has ? i : -1;
            skip[i] = (has && (wp[i].mbBad || found.count(&wp[i]))) ? 1 : 0;
            std::memcpy(&world[(size_t)i * 3], wp[i].mWorldPos, 12);
            mind[i] = wp[i].mfMinDistance;
            std::memcpy(&mpd[(size_t)i * 32], wp[i].mDescriptor, 32);
        }
        // vocabulary stand-in: node = leading descriptor bits
        std::vector<int> kn(F1.N), fn(F2.N);
        for (int i = 0; i < F1.N; i++) { kn[i] = 40 + (F1.mDescriptors.ptr(i)[0] & 0x1f); Fk.mFeatVec[(unsigned)kn[i]].push_back((unsigned)i); }
        for (int j = 0; j < F2.N; j++) { fn[j] = 40 + (F2.mDescriptors.ptr(j)[0] & 0x1f); Fc.mFeatVec[(unsigned)fn[j]].push_back((unsigned)j); }
        KeyFrame KF(Fk);
        ORBmatcher mb(0.7f, true);
        oracle::ORBmatcher omb(0.7f, true);
        std::vector<MapPoint*> vm;
        std::vector<int> ovm;
        oracle::Frame oKb = oK;
        for (int i = 0; i < F1.N; i++) if (oKb.mvpMapPoints[i] >= 0 && wp[i].mbBad) oKb.mvpMapPoints[i] = -1;   // isBad() -> skipped
        nBow = mb.SearchByBoW(&KF, Fc, vm);
        const int onBow = omb.SearchByBoW(oKb, kn.data(), oC, fn.data(), ovm);
        CHECK(nBow == onBow && nBow > 50, "SearchByBoW: %d vs %d", nBow, onBow);
        bool okb = vm.size() == ovm.size();
        for (size_t j = 0; okb && j < vm.size(); j++) okb = (vm[j] ? (int)(vm[j] - &wp[0]) : -1) == ovm[j];
        CHECK(okb, "SearchByBoW associations differ");
        // relocalisation-style projection search: the current frame already holds a few points
        for (int j = 0; j < Fc.N; j++) {
            Fc.mvpMapPoints[j] = nullptr; oC.mvpMapPoints[j] = -1; Fc.mvpMatchScore[j] = 999; oC.mvpMatchScore[j] = 999;
            if ((splitmix() % 11) == 0) { Fc.mvpMapPoints[j] = &wp[0]; oC.mvpMapPoints[j] = F1.N; }
        }
        Fc.fx = fx; Fc.fy = fy; Fc.cx = cx; Fc.cy = cy; oC.fx = fx; oC.fy = fy; oC.cx = cx; oC.cy = cy;
        ORBmatcher mk(0.9f, true);
        oracle::ORBmatcher omk(0.9f, true);
        nKf = mk.SearchByProjection(Fc, &KF, found, 10.0f, 100);
        const int onKf = omk.SearchByProjection(oC, oK, skip.data(), world.data(), mind.data(), mpd.data(), 10.0f, 100);
        CHECK(nKf == onKf && nKf > 50, "SearchByProjection(Current, KeyFrame): %d vs %d", nKf, onKf);
        bool okk = true;
        for (int j = 0; okk && j < Fc.N; j++) {
            int id = -1;
            if (Fc.mvpMapPoints[j]) { const int i = (int)(Fc.mvpMapPoints[j] - &wp[0]); id = (oC.mvpMapPoints[j] == F1.N) ? F1.N : i; }
            okk = id == oC.mvpMapPoints[j] && Fc.mvpMatchScore[j] == oC.mvpMatchScore[j];
        }
        CHECK(okk, "SearchByProjection(Current, KeyFrame) associations / scores differ");
    }
    printf("keyframe leg: %d BoW matches, %d keyframe-projection matches\n", nBow, nKf);

    // ---- the overloads of the LocalMapping / LoopClosing threads through the shim vs the oracle (oracle/orb_matcher_kf.cpp):
    //      SearchByBoW(KF, KF), SearchForTriangulation, SearchByProjection(KF, Scw, ..), SearchBySim3, Fuse(KF, Scw, ..), Fuse(KF, ..)
    int nBow2 = 0, nTri = 0, nSbp = 0, nSim3 = 0, nFuseS = 0, nFuse = 0;
    {
        const float fx = 458.0f, fy = 457.0f, cx = 320.0f, cy = 240.0f;
        const int N1 = F1.N, N2 = F2.N;
        sm_state = 9191;
        float sc[8]; sc[0] = 1.0f; for (int l = 1; l < 8; l++) sc[l] = sc[l - 1] * 1.2f;
        Frame Fa = F1, Fb = F2;
        for (Frame* f : {&Fa, &Fb}) { f->fx = fx; f->fy = fy; f->cx = cx; f->cy = cy; f->mFeatVec.clear(); }
        std::vector<int> nd1(N1), nd2(N2);
        for (int i = 0; i < N1; i++) { nd1[i] = 40 + (F1.mDescriptors.ptr(i)[0] & 0x1f); Fa.mFeatVec[(unsigned)nd1[i]].push_back((unsigned)i); }
        for (int j = 0; j < N2; j++) { nd2[j] = 40 + (F2.mDescriptors.ptr(j)[0] & 0x1f); Fb.mFeatVec[(unsigned)nd2[j]].push_back((unsigned)j); }
        KeyFrame KF1(Fa), KF2(Fb);
        const float eye[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
        std::memcpy(KF1.Tcw, eye, sizeof(eye)); std::memcpy(KF2.Tcw, eye, sizeof(eye));
        // map points: ids 0 .. N1-1 sit on KF1's keypoints and project (identity poses) onto the matching KF2 keypoints,
        // ids N1 .. N1+N2-1 sit on KF2's keypoints and project onto KF1's
        std::vector<MapPoint> pts((size_t)N1 + N2);
        std::vector<oracle::MapPointGeom> geom(pts.size());
        auto make = [&](int id, const gforb_cv::KeyPoint& kp, float du, float dv, const uint8_t* d) {
            MapPoint& p = pts[id];
            const float z = 2.0f + (float)(splitmix() % 600) * 0.01f;
            const float u = kp.pt.x + du, v = kp.pt.y + dv;
            p.mWorldPos[0] = (u - cx) / fx * z; p.mWorldPos[1] = (v - cy) / fy * z; p.mWorldPos[2] = z;
            const float d3 = std::sqrt(p.mWorldPos[0] * p.mWorldPos[0] + p.mWorldPos[1] * p.mWorldPos[1] + z * z);
            p.mfMinDistance = d3 / (sc[kp.octave] * (0.95f + 0.01f * (float)(splitmix() % 10)));
            p.mfMaxDistance = p.mfMinDistance * 3.58f * ((splitmix() % 25) == 0 ? 0.2f : 1.3f);
            p.mNormalVector[0] = 0; p.mNormalVector[1] = 0; p.mNormalVector[2] = ((splitmix() % 20) == 0) ? -1.0f : 1.0f;
            std::memcpy(p.mDescriptor, d, 32);
            if ((splitmix() % 5) == 0) p.mDescriptor[splitmix() % 32] ^= 0x22;
            p.mbBad = (splitmix() % 30) == 0;
        };
        for (int i = 0; i < N1; i++) make(i, k1[i], -3.0f, -1.0f, F1.mDescriptors.ptr(i));
        for (int j = 0; j < N2; j++) make(N1 + j, k2[j], 3.0f, 1.0f, F2.mDescriptors.ptr(j));
        for (size_t id = 0; id < pts.size(); id++) {
            std::memcpy(geom[id].world, pts[id].mWorldPos, 12); std::memcpy(geom[id].normal, pts[id].mNormalVector, 12);
            geom[id].minDist = pts[id].mfMinDistance; geom[id].maxDist = pts[id].mfMaxDistance;
            std::memcpy(geom[id].desc, pts[id].mDescriptor, 32);
        }
        for (int i = 0; i < N1; i++) { const bool has = (splitmix() % 2) != 0; KF1.mvpMapPoints[i] = has ? &pts[i] : nullptr; if (has) pts[i].AddObservation(&KF1, i); }
        for (int j = 0; j < N2; j++) { const bool has = (splitmix() % 2) != 0; KF2.mvpMapPoints[j] = has ? &pts[N1 + j] : nullptr; if (has) pts[N1 + j].AddObservation(&KF2, j); }
        auto idOf = [&](const MapPoint* p) { return p ? (int)(p - &pts[0]) : -1; };
        auto oracleKF = [&](const oracle::Frame& base, const KeyFrame& KF, bool dropBad) {
            oracle::Frame o = base;
            o.fx = fx; o.fy = fy; o.cx = cx; o.cy = cy;
            std::memcpy(o.mTcw, eye, sizeof(eye));
            for (int i = 0; i < KF.N; i++) {
                const MapPoint* p = KF.mvpMapPoints[i];
                o.mvpMapPoints[i] = (p && !(dropBad && p->isBad())) ? idOf(p) : -1;
            }
            return o;
        };
        auto badVec = [&]() { std::vector<uint8_t> b(pts.size()); for (size_t i = 0; i < pts.size(); i++) b[i] = pts[i].isBad() ? 1 : 0; return b; };

        {   // SearchByBoW(pKF1, pKF2, vpMatches12)
            ORBmatcher mb(0.8f, true);
            oracle::ORBmatcher omb(0.8f, true);
            std::vector<MapPoint*> vm;
            std::vector<int> ovm;
            nBow2 = mb.SearchByBoW(&KF1, &KF2, vm);
            const int on = omb.SearchByBoW_KeyFrames(oracleKF(oF1, KF1, true), nd1.data(), oracleKF(oF2, KF2, true), nd2.data(), ovm);
            bool ok = vm.size() == ovm.size();
            for (size_t i = 0; ok && i < vm.size(); i++) ok = idOf(vm[i]) == ovm[i];
            CHECK(nBow2 == on && nBow2 > 20 && ok, "SearchByBoW(KF, KF): %d vs %d", nBow2, on);
        }
        {   // SearchForTriangulation: frame 2 is frame 1 shifted by (3, 1): epipolar lines y2 = y1 (one row off passes the gate)
            gforb_cv::Mat F12; F12.create(3, 3, 4);
            const float Fv[9] = {0, 1e-5f, 0, -1e-5f, 0, -1, 2e-5f, 1, 0.25f};
            for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) F12.at<float>(r, c) = Fv[r * 3 + c];
            ORBmatcher mt(0.6f, true);
            oracle::ORBmatcher omt(0.6f, true);
            std::vector<gforb_cv::KeyPoint> mk1, mk2;
            std::vector<std::pair<size_t, size_t>> pairs;
            std::vector<int> ovm;
            nTri = mt.SearchForTriangulation(&KF1, &KF2, F12, mk1, mk2, pairs);
            const int on = omt.SearchForTriangulation(oracleKF(oF1, KF1, false), nd1.data(), oracleKF(oF2, KF2, false), nd2.data(), Fv, ovm);
            size_t k = 0; bool ok = true;
            for (int i = 0; i < N1; i++) if (ovm[i] >= 0) { ok = ok && k < pairs.size() && pairs[k].first == (size_t)i && pairs[k].second == (size_t)ovm[i]; k++; }
            ok = ok && k == pairs.size() && mk1.size() == k && mk2.size() == k;
            CHECK(nTri == on && nTri > 20 && ok, "SearchForTriangulation: %d vs %d", nTri, on);
        }
        gforb_cv::Mat Scw; Scw.create(4, 4, 4);
        float Sv[16] = {0};
        {
            const float s = 1.3f, a = 0.001f;
            const float Rz[9] = {std::cos(a), -std::sin(a), 0, std::sin(a), std::cos(a), 0, 0, 0, 1};
            for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) Sv[r * 4 + c] = s * Rz[r * 3 + c]; }
            Sv[3] = s * 0.002f; Sv[7] = s * -0.001f; Sv[11] = s * 0.003f; Sv[15] = 1.0f;
            for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) Scw.at<float>(r, c) = Sv[r * 4 + c];
        }
        {   // SearchByProjection(pKF2, Scw, vpPoints, vpMatched, th): KF1's points (with repeats) into KF2
            std::vector<MapPoint*> list, vpMatched(N2, nullptr);
            std::vector<int> olist, omatched(N2, -1);
            for (int i = 0; i < N1; i++) { list.push_back(&pts[i]); if ((splitmix() % 9) == 0) list.push_back(&pts[splitmix() % N1]); }
            for (int j = 0; j < N2; j++) if ((splitmix() % 8) == 0) vpMatched[j] = &pts[splitmix() % N1];
            for (MapPoint* p : list) olist.push_back(idOf(p));
            for (int j = 0; j < N2; j++) omatched[j] = idOf(vpMatched[j]);
            ORBmatcher ms(0.6f, true);
            oracle::ORBmatcher oms(0.6f, true);
            nSbp = ms.SearchByProjection(&KF2, Scw, list, vpMatched, 10);
            const std::vector<uint8_t> bad = badVec();
            const int on = oms.SearchByProjection(oracleKF(oF2, KF2, false), Sv, olist, geom.data(), bad.data(), omatched, 10);
            bool ok = true;
            for (int j = 0; j < N2; j++) ok = ok && idOf(vpMatched[j]) == omatched[j];
            CHECK(nSbp == on && nSbp > 50 && ok, "SearchByProjection(KF, Scw): %d vs %d", nSbp, on);
        }
        {   // SearchBySim3: a similarity close to the identity between the two (identity-pose) keyframes
            gforb_cv::Mat R12, t12; R12.create(3, 3, 4); t12.create(3, 1, 4);
            const float a = 0.0015f, s12 = 1.004f;
            const float Rv[9] = {std::cos(a), -std::sin(a), 0, std::sin(a), std::cos(a), 0, 0, 0, 1}, tv[3] = {0.003f, -0.002f, 0.004f};
            for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) R12.at<float>(r, c) = Rv[r * 3 + c]; t12.at<float>(r, 0) = tv[r]; }
            std::vector<MapPoint*> vm(N1, nullptr);
            std::vector<int> ovm(N1, -1), idx2(N1, -1);
            for (int i = 0; i < N1; i++)
                if (KF1.mvpMapPoints[i] && (splitmix() % 12) == 0) {
                    const int j = (int)(splitmix() % N2);
                    if (KF2.mvpMapPoints[j]) { vm[i] = KF2.mvpMapPoints[j]; ovm[i] = idOf(vm[i]); idx2[i] = vm[i]->GetIndexInKeyFrame(&KF2); }
                }
            ORBmatcher m3(0.6f, true);
            oracle::ORBmatcher om3(0.6f, true);
            nSim3 = m3.SearchBySim3(&KF1, &KF2, vm, s12, R12, t12, 7.5f);
            const std::vector<uint8_t> bad = badVec();
            const int on = om3.SearchBySim3(oracleKF(oF1, KF1, false), oracleKF(oF2, KF2, false), ovm, idx2.data(), geom.data(), bad.data(), s12, Rv, tv, 7.5f);
            bool ok = true;
            for (int i = 0; i < N1; i++) ok = ok && idOf(vm[i]) == ovm[i];
            CHECK(nSim3 == on && nSim3 > 30 && ok, "SearchBySim3: %d vs %d", nSim3, on);
        }
        // the Fuse overloads change the map: the oracle's model of the keyframe (slots, bad flags, observation index) is taken
        // from the shim's state before each call and compared after it
        auto fuseCase = [&](KeyFrame& KF, const oracle::Frame& obase, int firstId, int nIds, bool sim3, int& nOut) {
            std::vector<MapPoint*> list;
            for (int i = 0; i < nIds; i++) {
                list.push_back(&pts[firstId + i]);
                if ((splitmix() % 10) == 0) list.push_back(&pts[firstId + (int)(splitmix() % nIds)]);     // repeated entries
                if (!sim3 && (splitmix() % 25) == 0) list.push_back(nullptr);
            }
            std::vector<int> olist, slot(KF.N), inKF(pts.size(), -1), obi, obd;
            for (MapPoint* p : list) olist.push_back(idOf(p));
            for (int j = 0; j < KF.N; j++) slot[j] = idOf(KF.mvpMapPoints[j]);
            for (size_t id = 0; id < pts.size(); id++) inKF[id] = pts[id].GetIndexInKeyFrame(&KF);
            std::vector<uint8_t> bad = badVec();
            ORBmatcher mf(0.6f, true);
            oracle::ORBmatcher omf(0.6f, true);
            const oracle::Frame okf = oracleKF(obase, KF, false);
            int on;
            if (sim3) { nOut = mf.Fuse(&KF, Scw, list, 4.0f); on = omf.Fuse(okf, Sv, olist, geom.data(), 4.0f, slot, bad, inKF, obi, obd); }
            else { nOut = mf.Fuse(&KF, list, 3.0f); on = omf.Fuse(okf, olist, geom.data(), 3.0f, slot, bad, inKF, obi, obd); }
            bool ok = true;
            for (int j = 0; j < KF.N; j++) ok = ok && idOf(KF.mvpMapPoints[j]) == slot[j];
            for (size_t id = 0; id < pts.size(); id++) ok = ok && (pts[id].isBad() ? 1 : 0) == bad[id] && pts[id].GetIndexInKeyFrame(&KF) == inKF[id];
            CHECK(nOut == on && nOut > 30 && ok, "Fuse(%s): %d vs %d fused, map state %s", sim3 ? "KF, Scw" : "KF", nOut, on, ok ? "equal" : "differs");
        };
        fuseCase(KF1, oF1, N1, N2, true, nFuseS);          // KF2's points fused into KF1 through Scw
        fuseCase(KF2, oF2, 0, N1, false, nFuse);           // KF1's points fused into KF2 with its own pose
    }
    printf("mapping / loop-closing leg: %d BoW(KF, KF), %d triangulation, %d projection(Scw), %d Sim3, %d + %d fused\n", nBow2, nTri, nSbp,
           nSim3, nFuseS, nFuse);

    printf("TrackPreviousFrame leg: %d window-projection matches, %d pose inliers\n", nPrev, nInl);
    printf("shim test: %zu/%zu keypoints, %d init matches, %d projection matches, %zu good features, %d active matches: %s\n", k1.size(),
           k2.size(), nm, np, good.size(), nActive, g_fail ? "FAILED" : "OK");
    return g_fail ? 1 : 0;
}
