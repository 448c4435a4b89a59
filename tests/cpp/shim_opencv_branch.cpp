// Compile check of include/gforb/orb_slam.hpp with -DGFORB_WITH_OPENCV against the OpenCV interface stand-in
// (tests/cpp/opencv_stub): instantiates every class and names every member a maintainer's call sites use.
#define GFORB_WITH_OPENCV 1
#include "gforb/orb_slam.hpp"

using namespace ORB_SLAM;

int touch_everything(bool run) {
    if (!run) return 0;
    std::vector<cv::KeyPoint> keys;
    cv::Mat image(480, 752, CV_8UC1), mask, desc;
    ORBextractor ex(1000, 1.2f, 8, ORBextractor::FAST_SCORE, 20);
    ex(image, mask, keys, desc);
    Frame F1(keys, desc, 752, 480), F2(keys, desc, 752, 480);
    ORBmatcher m(0.9f, true);
    std::vector<cv::Point2f> prev(F1.N);
    std::vector<int> m12;
    int n = m.SearchForInitialization(F1, F2, prev, m12, 100);
    std::vector<MapPoint*> mps, out;
    n += m.SearchByProjection(F1, mps, 3.0f);
    n += m.SearchByProjection(F2, F1, 15.0f);
    n += m.WindowSearch(F1, F2, 200, out, 0, 7);
    n += m.SearchByProjection(F1, F2, 15, out);
    n += ORBmatcher::DescriptorDistance(desc.row(0), desc.row(0));
    KeyFrame K1(F1), K2(F2);
    std::set<MapPoint*> found;
    n += m.SearchByBoW(&K1, F2, out);
    n += m.SearchByProjection(F2, &K1, found, 10.0f, 100);
    n += m.SearchByBoW(&K1, &K2, out);
    cv::Mat F12(3, 3, CV_32F), Scw(4, 4, CV_32F), R12(3, 3, CV_32F), t12(3, 1, CV_32F);
    std::vector<cv::KeyPoint> a, b;
    std::vector<std::pair<size_t, size_t>> pairs;
    n += m.SearchForTriangulation(&K1, &K2, F12, a, b, pairs);
    n += m.Fuse(&K1, mps, 3.0f);
    n += m.Fuse(&K1, Scw, mps, 4.0f);
    n += m.SearchByProjection(&K1, Scw, mps, out, 10);
    n += m.SearchBySim3(&K1, &K2, out, 1.0f, R12, t12, 7.5f);
    Observability obs(458.0, 457.0, 480, 752, 367.0, 248.0, 0.0, 0.0);
    cv::Mat T(4, 4, CV_32F);
    obs.updatePWLSVec(0.0, T, 0.05, T);
    obs.predictPWLSVec(0.05, 1);
    std::vector<GoodPoint> good;
    obs.pFrame = &F1;
    n += obs.runMatrixBuilding(ORB_SLAM::FRAME_INFO_MATRIX, 1.0, true) ? 1 : 0;
    n += obs.setSelction_Number(100, 2, 1.0, 0.1, &F1, &good) ? 1 : 0;
    n += Optimizer::PoseOptimization(&F1);
    return n;
}

int main() { return touch_everything(false); }
