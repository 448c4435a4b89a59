// TEST INFRASTRUCTURE ONLY (see oracle.hpp).  ORBmatcher overloads of the LocalMapping / LoopClosing threads,
// restated from /root/reference/src/ORBmatcher.cc.  KeyFrames are oracle::Frame objects (same grid, same keypoint arrays); map
// points are integer ids.  The reference holds no vectors for these functions: parity unpinned w.r.t. its tests; the cv::Mat
// arithmetic is pinned against cv2 4.13 where Python can reach it (tests/golden/cv413_gemm_small.npz).
#include <algorithm>
#include <climits>
#include <cmath>
#include <map>
#include <set>
#include "oracle.hpp"

namespace oracle {

// src/ORBmatcher.cc:866-871 / :1718-1723
Sim3Rt decomposeSim3(const float* Scw) {
    Sim3Rt o;
    // scw = sqrt(sRcw.row(0).dot(sRcw.row(0))): Mat::dot on CV_32F accumulates double products; the double root narrows to float
    double d = 0;
    for (int k = 0; k < 3; k++) d += (double)Scw[k] * (double)Scw[k];
    const float scw = std::sqrt(d);
    // sRcw/scw and Scw.col(3)/scw: operator/(Mat, double) = MatExpr scale by 1./s -> convertTo(CV_32F, alpha)
    const double alpha = 1. / (double)scw;
    for (int r = 0; r < 3; r++) {
        for (int c = 0; c < 3; c++) o.R[r * 3 + c] = cvScale32f(Scw[r * 4 + c], alpha);
        o.t[r] = cvScale32f(Scw[r * 4 + 3], alpha);
    }
    cvNegRtTimesT(o.R, 3, o.t, 1, o.Ow);
    return o;
}

// src/ORBmatcher.cc:705-722; sigma2kp2 = pKF2->GetSigma2(kp2.octave)
bool ORBmatcher::CheckDistEpipolarLine(const KeyPoint& kp1, const KeyPoint& kp2, const float* F12, float sigma2kp2) {
    const float a = kp1.x * F12[0 * 3 + 0] + kp1.y * F12[1 * 3 + 0] + F12[2 * 3 + 0];
    const float b = kp1.x * F12[0 * 3 + 1] + kp1.y * F12[1 * 3 + 1] + F12[2 * 3 + 1];
    const float c = kp1.x * F12[0 * 3 + 2] + kp1.y * F12[1 * 3 + 2] + F12[2 * 3 + 2];
    const float num = a * kp2.x + b * kp2.y + c;
    const float den = a * a + b * b;
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * sigma2kp2;
}

typedef std::map<int, std::vector<unsigned>> FeatVec;      // DBoW2::FeatureVector (Thirdparty/DBoW2/DBoW2/FeatureVector.h)
static FeatVec featVec(const Frame& F, const int* node) {
    FeatVec v;
    for (int i = 0; i < F.N; i++) if (node[i] >= 0) v[node[i]].push_back((unsigned)i);
    return v;
}

static void pruneByRotation(const std::vector<int>* rotHist, std::vector<int>& matches12, int& nmatches) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    ORBmatcher::ComputeThreeMaxima(rotHist, ORBmatcher::HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < ORBmatcher::HISTO_LENGTH; i++) {
        if (i == ind1 || i == ind2 || i == ind3) continue;
        for (int idx : rotHist[i]) { matches12[idx] = -1; nmatches--; }
    }
}

static int rotBin(float a1, float a2) {
    const float factor = 1.0f / ORBmatcher::HISTO_LENGTH;
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = std::round(rot * factor);
    if (bin == ORBmatcher::HISTO_LENGTH) bin = 0;
    return bin;
}

// src/ORBmatcher.cc:1289-1424.  KF*.mvpMapPoints: id or -1 (NULL / isBad()).
int ORBmatcher::SearchByBoW_KeyFrames(const Frame& KF1, const int* node1, const Frame& KF2, const int* node2, std::vector<int>& vpMatches12) {
    const std::vector<int>& vpMapPoints1 = KF1.mvpMapPoints;
    const std::vector<int>& vpMapPoints2 = KF2.mvpMapPoints;
    const FeatVec vFeatVec1 = featVec(KF1, node1), vFeatVec2 = featVec(KF2, node2);
    vpMatches12.assign(vpMapPoints1.size(), -1);
    std::vector<bool> vbMatched2(vpMapPoints2.size(), false);
    std::vector<int> rotHist[HISTO_LENGTH];
    int nmatches = 0;
    auto f1it = vFeatVec1.begin(), f1end = vFeatVec1.end();
    auto f2it = vFeatVec2.begin(), f2end = vFeatVec2.end();
    while (f1it != f1end && f2it != f2end) {
        if (f1it->first == f2it->first) {
            for (size_t i1 = 0; i1 < f1it->second.size(); i1++) {
                const size_t idx1 = f1it->second[i1];
                if (vpMapPoints1[idx1] < 0) continue;
                const uint8_t* d1 = &KF1.mDescriptors[idx1 * 32];
                int bestDist1 = INT_MAX, bestIdx2 = -1, bestDist2 = INT_MAX;
                for (size_t i2 = 0; i2 < f2it->second.size(); i2++) {
                    const size_t idx2 = f2it->second[i2];
                    if (vbMatched2[idx2] || vpMapPoints2[idx2] < 0) continue;
                    const int dist = DescriptorDistance(d1, &KF2.mDescriptors[idx2 * 32]);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = (int)idx2; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (bestDist1 < TH_LOW) {
                    if ((float)bestDist1 < mfNNratio * (float)bestDist2) {
                        vpMatches12[idx1] = vpMapPoints2[bestIdx2];
                        vbMatched2[bestIdx2] = true;
                        if (mbCheckOrientation) rotHist[rotBin(KF1.mvKeysUn[idx1].angle, KF2.mvKeysUn[bestIdx2].angle)].push_back((int)idx1);
                        nmatches++;
                    }
                }
            }
            f1it++;
            f2it++;
        } else if (f1it->first < f2it->first) {
            f1it = vFeatVec1.lower_bound(f2it->first);
        } else {
            f2it = vFeatVec2.lower_bound(f1it->first);
        }
    }
    if (mbCheckOrientation) pruneByRotation(rotHist, vpMatches12, nmatches);
    return nmatches;
}

// src/ORBmatcher.cc:1426-1588.  vMatches12[idx1] = idx2 or -1; the reference's vMatchedKeys1/2 and vMatchedPairs are the
// pairs (i, vMatches12[i]) in ascending i (:1577-1585).
int ORBmatcher::SearchForTriangulation(const Frame& KF1, const int* node1, const Frame& KF2, const int* node2, const float* F12,
                                       std::vector<int>& vMatches12) {
    const std::vector<int>& vpMapPoints1 = KF1.mvpMapPoints;
    const std::vector<int>& vpMapPoints2 = KF2.mvpMapPoints;
    const FeatVec vFeatVec1 = featVec(KF1, node1), vFeatVec2 = featVec(KF2, node2);
    int nmatches = 0;
    std::vector<bool> vbMatched2(KF2.mvKeysUn.size(), false);
    vMatches12.assign(KF1.mvKeysUn.size(), -1);
    std::vector<int> rotHist[HISTO_LENGTH];
    auto f1it = vFeatVec1.begin(), f1end = vFeatVec1.end();
    auto f2it = vFeatVec2.begin(), f2end = vFeatVec2.end();
    while (f1it != f1end && f2it != f2end) {
        if (f1it->first == f2it->first) {
            for (size_t i1 = 0; i1 < f1it->second.size(); i1++) {
                const size_t idx1 = f1it->second[i1];
                if (vpMapPoints1[idx1] >= 0) continue;                       // there is already a MapPoint
                const KeyPoint& kp1 = KF1.mvKeysUn[idx1];
                const uint8_t* d1 = &KF1.mDescriptors[idx1 * 32];
                std::vector<std::pair<int, size_t>> vDistIndex;
                for (size_t i2 = 0; i2 < f2it->second.size(); i2++) {
                    const size_t idx2 = f2it->second[i2];
                    if (vbMatched2[idx2] || vpMapPoints2[idx2] >= 0) continue;
                    const int dist = DescriptorDistance(d1, &KF2.mDescriptors[idx2 * 32]);
                    if (dist > TH_LOW) continue;
                    vDistIndex.push_back(std::make_pair(dist, idx2));
                }
                if (vDistIndex.empty()) continue;
                std::sort(vDistIndex.begin(), vDistIndex.end());
                const int BestDist = vDistIndex.front().first;
                const int DistTh = std::round(2 * BestDist);
                for (size_t id = 0; id < vDistIndex.size(); id++) {
                    if (vDistIndex[id].first > DistTh) break;
                    const int currentIdx2 = (int)vDistIndex[id].second;
                    const KeyPoint& kp2 = KF2.mvKeysUn[currentIdx2];
                    if (CheckDistEpipolarLine(kp1, kp2, F12, KF2.mvLevelSigma2[kp2.octave])) {
                        vbMatched2[currentIdx2] = true;
                        vMatches12[idx1] = currentIdx2;
                        nmatches++;
                        if (mbCheckOrientation) rotHist[rotBin(kp1.angle, kp2.angle)].push_back((int)idx1);
                        break;
                    }
                }
            }
            f1it++;
            f2it++;
        } else if (f1it->first < f2it->first) {
            f1it = vFeatVec1.lower_bound(f2it->first);
        } else {
            f2it = vFeatVec2.lower_bound(f1it->first);
        }
    }
    if (mbCheckOrientation) pruneByRotation(rotHist, vMatches12, nmatches);
    return nmatches;
}

// The projection block the two Fuse overloads and SearchByProjection(KF, Scw, ...) share (:1617-1687, :1743-1816, :889-965):
// returns false where the reference `continue`s before the descriptor loop; otherwise bestIdx / bestDist of the loop.
// claimed: vpMatched of SearchByProjection (ids, skipped when >= 0) or nullptr (Fuse looks at every keypoint).
static bool projectAndMatch(const Frame& KF, const float* Rcw, int rstep, const float* tcw, int tstep, const float* Ow,
                            const MapPointGeom& g, float th, const std::vector<int>* claimed, int& bestIdx, int& bestDist) {
    const int nMaxLevel = (int)KF.mvScaleFactors.size() - 1;
    float p3Dc[3];
    cvGemmRxPlusT(Rcw, rstep, g.world, tcw, tstep, p3Dc);
    if (p3Dc[2] < 0.0f) return false;
    const float invz = 1.0 / p3Dc[2];                       // 1/z (float) and 1.0/z (double, narrowed) round alike
    const float x = p3Dc[0] * invz, y = p3Dc[1] * invz;
    const float u = KF.fx * x + KF.cx, v = KF.fy * y + KF.cy;
    if (!(u >= KF.mnMinX && u < KF.mnMaxX && v >= KF.mnMinY && v < KF.mnMaxY)) return false;      // KeyFrame::IsInImage (:654)
    const float PO[3] = {g.world[0] - Ow[0], g.world[1] - Ow[1], g.world[2] - Ow[2]};
    const float dist3D = std::sqrt((double)PO[0] * PO[0] + (double)PO[1] * PO[1] + (double)PO[2] * PO[2]);   // cv::norm
    if (dist3D < g.minDist || dist3D > g.maxDist) return false;
    const double dot = (double)PO[0] * g.normal[0] + (double)PO[1] * g.normal[1] + (double)PO[2] * g.normal[2];   // Mat::dot
    if (dot < 0.5 * dist3D) return false;
    const float ratio = dist3D / g.minDist;
    auto it = std::lower_bound(KF.mvScaleFactors.begin(), KF.mvScaleFactors.end(), ratio);
    const int nPredictedLevel = std::min((int)(it - KF.mvScaleFactors.begin()), nMaxLevel);
    const float radius = th * KF.mvScaleFactors[nPredictedLevel];
    const std::vector<size_t> vIndices = KF.GetFeaturesInArea(u, v, radius);
    if (vIndices.empty()) return false;
    bestDist = INT_MAX; bestIdx = -1;
    for (size_t idx : vIndices) {
        if (claimed && (*claimed)[idx] >= 0) continue;
        const int kpLevel = KF.mvKeysUn[idx].octave;
        if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
        const int dist = ORBmatcher::DescriptorDistance(g.desc, &KF.mDescriptors[idx * 32]);
        if (dist < bestDist) { bestDist = dist; bestIdx = (int)idx; }
    }
    return true;
}

// src/ORBmatcher.cc:1590-1708.  KF.mTcw = pose, Ow = pKF->GetCameraCenter() = -Rcw.t()*tcw (src/KeyFrame.cc:73,82).
// MapPoint::Replace / AddObservation / KeyFrame::AddMapPoint (src/MapPoint.cc:136-170) are modelled on this keyframe only.
int ORBmatcher::Fuse(const Frame& KF, const std::vector<int>& points, const MapPointGeom* geom, float th, std::vector<int>& kfSlot,
                     std::vector<uint8_t>& bad, std::vector<int>& idxInKF, std::vector<int>& bestIdxOut, std::vector<int>& bestDistOut) {
    const float* T = KF.mTcw;
    float Ow[3];
    cvNegRtTimesT(T, 4, T + 3, 4, Ow);
    bestIdxOut.assign(points.size(), -1); bestDistOut.assign(points.size(), INT_MAX);
    int nFused = 0;
    for (size_t i = 0; i < points.size(); i++) {
        const int pMP = points[i];
        if (pMP < 0) continue;
        if (bad[pMP] || idxInKF[pMP] >= 0) continue;                         // isBad() || IsInKeyFrame(pKF)
        int bestIdx, bestDist;
        if (!projectAndMatch(KF, T, 4, T + 3, 4, Ow, geom[pMP], th, nullptr, bestIdx, bestDist)) continue;
        bestIdxOut[i] = bestIdx; bestDistOut[i] = bestDist;
        if (bestDist <= TH_LOW) {
            const int pMPinKF = kfSlot[bestIdx];
            if (pMPinKF >= 0) {
                if (!bad[pMPinKF]) bad[pMP] = 1;                             // pMP->Replace(pMPinKF): pMP holds no observation in pKF
            } else {
                idxInKF[pMP] = bestIdx;                                      // pMP->AddObservation(pKF, bestIdx)
                kfSlot[bestIdx] = pMP;                                       // pKF->AddMapPoint(pMP, bestIdx)
            }
            nFused++;
        }
    }
    return nFused;
}

// src/ORBmatcher.cc:1710-1839
int ORBmatcher::Fuse(const Frame& KF, const float* Scw, const std::vector<int>& points, const MapPointGeom* geom, float th,
                     std::vector<int>& kfSlot, std::vector<uint8_t>& bad, std::vector<int>& idxInKF, std::vector<int>& bestIdxOut,
                     std::vector<int>& bestDistOut) {
    const Sim3Rt S = decomposeSim3(Scw);
    std::set<int> spAlreadyFound;                                            // pKF->GetMapPoints() (src/KeyFrame.cc:238-251)
    for (int id : kfSlot) if (id >= 0 && !bad[id]) spAlreadyFound.insert(id);
    bestIdxOut.assign(points.size(), -1); bestDistOut.assign(points.size(), INT_MAX);
    int nFused = 0;
    for (size_t iMP = 0; iMP < points.size(); iMP++) {
        const int pMP = points[iMP];
        if (bad[pMP] || spAlreadyFound.count(pMP)) continue;
        int bestIdx, bestDist;
        if (!projectAndMatch(KF, S.R, 3, S.t, 1, S.Ow, geom[pMP], th, nullptr, bestIdx, bestDist)) continue;
        bestIdxOut[iMP] = bestIdx; bestDistOut[iMP] = bestDist;
        if (bestDist <= TH_LOW) {
            const int pMPinKF = kfSlot[bestIdx];
            if (pMPinKF >= 0) {
                if (!bad[pMPinKF] && pMPinKF != pMP) {                       // pMPinKF->Replace(pMP) (same id: returns at once)
                    bad[pMPinKF] = 1;
                    const int at = idxInKF[pMPinKF];                         // its observation in pKF
                    idxInKF[pMPinKF] = -1;
                    if (at >= 0) {
                        if (idxInKF[pMP] < 0) { kfSlot[at] = pMP; idxInKF[pMP] = at; }   // ReplaceMapPointMatch + AddObservation
                        else kfSlot[at] = -1;                                             // EraseMapPointMatch
                    }
                }
            } else {
                idxInKF[pMP] = bestIdx;
                kfSlot[bestIdx] = pMP;
            }
            nFused++;
        }
    }
    return nFused;
}

// src/ORBmatcher.cc:855-977
int ORBmatcher::SearchByProjection(const Frame& KF, const float* Scw, const std::vector<int>& points, const MapPointGeom* geom,
                                   const uint8_t* bad, std::vector<int>& vpMatched, int th) {
    const Sim3Rt S = decomposeSim3(Scw);
    std::set<int> spAlreadyFound(vpMatched.begin(), vpMatched.end());
    spAlreadyFound.erase(-1);
    int nmatches = 0;
    for (size_t iMP = 0; iMP < points.size(); iMP++) {
        const int pMP = points[iMP];
        if (bad[pMP] || spAlreadyFound.count(pMP)) continue;
        int bestIdx, bestDist;
        if (!projectAndMatch(KF, S.R, 3, S.t, 1, S.Ow, geom[pMP], (float)th, &vpMatched, bestIdx, bestDist)) continue;
        if (bestDist <= TH_LOW) {
            vpMatched[bestIdx] = pMP;
            nmatches++;
        }
    }
    return nmatches;
}

// one direction of SearchBySim3 (:1893-1973 / :1976-2058): map points of KFa, camera A from world (Ta), then the similarity
// (sRba, tba) into camera B, searched in KFb.  Returns vnMatch (per KFa keypoint).
static std::vector<int> sim3Direction(const Frame& KFa, const Frame& KFb, const std::vector<bool>& vbAlreadyMatched,
                                      const MapPointGeom* geom, const uint8_t* bad, const float* sRba, const float* tba, float th) {
    const int nMaxLevel = (int)KFb.mvScaleFactors.size() - 1;
    std::vector<int> vnMatch(KFa.mvpMapPoints.size(), -1);
    for (size_t i = 0; i < KFa.mvpMapPoints.size(); i++) {
        const int pMP = KFa.mvpMapPoints[i];
        if (pMP < 0 || vbAlreadyMatched[i]) continue;
        if (bad[pMP]) continue;
        const MapPointGeom& g = geom[pMP];
        float p3Dca[3], p3Dcb[3];
        cvGemmRxPlusT(KFa.mTcw, 4, g.world, KFa.mTcw + 3, 4, p3Dca);        // R1w*p3Dw + t1w
        cvGemmRxPlusT(sRba, 3, p3Dca, tba, 1, p3Dcb);                       // sR21*p3Dc1 + t21
        if (p3Dcb[2] < 0.0) continue;
        const float invz = 1.0 / p3Dcb[2];
        const float x = p3Dcb[0] * invz, y = p3Dcb[1] * invz;
        const float u = KFa.fx * x + KFa.cx, v = KFa.fy * y + KFa.cy;       // both directions use pKF1's intrinsics (:1844-1847)
        if (!(u >= KFb.mnMinX && u < KFb.mnMaxX && v >= KFb.mnMinY && v < KFb.mnMaxY)) continue;
        const float dist3D = std::sqrt((double)p3Dcb[0] * p3Dcb[0] + (double)p3Dcb[1] * p3Dcb[1] + (double)p3Dcb[2] * p3Dcb[2]);
        if (dist3D < g.minDist || dist3D > g.maxDist) continue;
        const float ratio = dist3D / g.minDist;
        auto it = std::lower_bound(KFb.mvScaleFactors.begin(), KFb.mvScaleFactors.end(), ratio);
        const int nPredictedLevel = std::min((int)(it - KFb.mvScaleFactors.begin()), nMaxLevel);
        const float radius = th * KFb.mvScaleFactors[nPredictedLevel];
        const std::vector<size_t> vIndices = KFb.GetFeaturesInArea(u, v, radius);
        if (vIndices.empty()) continue;
        int bestDist = INT_MAX, bestIdx = -1;
        for (size_t idx : vIndices) {
            const int oct = KFb.mvKeysUn[idx].octave;
            if (oct < nPredictedLevel - 1 || oct > nPredictedLevel) continue;
            const int dist = ORBmatcher::DescriptorDistance(g.desc, &KFb.mDescriptors[idx * 32]);
            if (dist < bestDist) { bestDist = dist; bestIdx = (int)idx; }
        }
        if (bestDist <= ORBmatcher::TH_HIGH) vnMatch[i] = bestIdx;
    }
    return vnMatch;
}

// src/ORBmatcher.cc:1841-2079.  KF1.fx.. = pKF1's intrinsics (used for both projections, as in the reference).
int ORBmatcher::SearchBySim3(const Frame& KF1, const Frame& KF2, std::vector<int>& vpMatches12, const int* idx2InKF2,
                             const MapPointGeom* geom, const uint8_t* bad, float s12, const float* R12, const float* t12, float th) {
    // cv::Mat sR12 = s12*R12; sR21 = (1.0/s12)*R12.t(); t21 = -sR21*t12 (:1858-1860)
    float sR12[9], sR21[9], t21[3];
    for (int i = 0; i < 9; i++) sR12[i] = cvScale32f(R12[i], (double)s12);
    const double inv = 1.0 / s12;
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) sR21[r * 3 + c] = cvScale32f(R12[c * 3 + r], inv);
    for (int r = 0; r < 3; r++) {                            // gemm(sR21, t12, -1): small-matrix path, c = 0
        const float t0 = sR21[r * 3 + 0] * t12[0] + sR21[r * 3 + 1] * t12[1] + sR21[r * 3 + 2] * t12[2];
        t21[r] = (float)((double)t0 * -1.0 + 0.0);
    }
    const int N1 = (int)KF1.mvpMapPoints.size(), N2 = (int)KF2.mvpMapPoints.size();
    std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
    for (int i = 0; i < N1; i++) {
        if (vpMatches12[i] >= 0) {
            vbAlreadyMatched1[i] = true;
            const int idx2 = idx2InKF2[i];
            if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
        }
    }
    // the second direction projects with pKF1's fx, fy, cx, cy as well and tests pKF1->IsInImage (:1998-2002)
    Frame KF2cam = KF2;
    KF2cam.fx = KF1.fx; KF2cam.fy = KF1.fy; KF2cam.cx = KF1.cx; KF2cam.cy = KF1.cy;
    const std::vector<int> vnMatch1 = sim3Direction(KF1, KF2, vbAlreadyMatched1, geom, bad, sR21, t21, th);
    const std::vector<int> vnMatch2 = sim3Direction(KF2cam, KF1, vbAlreadyMatched2, geom, bad, sR12, t12, th);
    int nFound = 0;
    for (int i1 = 0; i1 < N1; i1++) {
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0) {
            const int idx1 = vnMatch2[idx2];
            if (idx1 == i1) {
                vpMatches12[i1] = KF2.mvpMapPoints[idx2];
                nFound++;
            }
        }
    }
    return nFound;
}

}  // namespace oracle
