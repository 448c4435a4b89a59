// TEST INFRASTRUCTURE ONLY (see oracle.hpp).  CPU restatement of the per-frame-tracking parts of
// ORB_SLAM::Frame (src/Frame.cc:118-131, 300-377) and ORB_SLAM::ORBmatcher (src/ORBmatcher.cc:384-465, 696-702,
// 979-1086, 1172-1287, 2081-2202, 2338-2400; include/ORBmatcher.h:71-145).
// MapPoint* / NULL is flattened to an int id / -1.
#include "oracle.hpp"
#include <set>
#include <algorithm>
#include <map>
#include <climits>
#include <cmath>
#include <cstring>

namespace oracle {

// src/Frame.cc:55-134 (the parts that matter to matching; no undistortion: mDistCoef(0)==0 -> mvKeysUn=mvKeys)
void Frame::init(const KeyPoint* kps, const uint8_t* desc, int n, int imgW, int imgH, int nlevels, float scaleFactor) {
    N = n;
    mvKeysUn.assign(kps, kps + n);
    mDescriptors.assign(desc, desc + (size_t)n * 32);
    mvpMapPoints.assign(n, -1);
    mvpMatchScore.assign(n, 999);
    mvbOutlier.assign(n, 0);
    mnMinX = 0; mnMaxX = imgW; mnMinY = 0; mnMaxY = imgH;              // src/Frame.cc:489-493
    mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(mnMaxX - mnMinX);
    mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(mnMaxY - mnMinY);
    mvScaleFactors.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvScaleFactors[0] = 1.0f; mvLevelSigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        mvScaleFactors[i] = mvScaleFactors[i - 1] * scaleFactor;
        mvLevelSigma2[i] = mvScaleFactors[i] * mvScaleFactors[i];
    }
    for (int i = 0; i < FRAME_GRID_COLS; i++)
        for (int j = 0; j < FRAME_GRID_ROWS; j++) mGrid[i][j].clear();
    for (size_t i = 0; i < mvKeysUn.size(); i++) {
        int gx, gy;
        if (PosInGrid(mvKeysUn[i], gx, gy)) mGrid[gx][gy].push_back(i);
    }
}

// src/Frame.cc:367-377
bool Frame::PosInGrid(const KeyPoint& kp, int& posX, int& posY) const {
    posX = std::round((kp.x - mnMinX) * mfGridElementWidthInv);
    posY = std::round((kp.y - mnMinY) * mfGridElementHeightInv);
    if (posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS) return false;
    return true;
}

// src/Frame.cc:300-365
std::vector<size_t> Frame::GetFeaturesInArea(float x, float y, float r, int minLevel, int maxLevel) const {
    std::vector<size_t> vIndices;
    int nMinCellX = std::floor((x - mnMinX - r) * mfGridElementWidthInv);
    nMinCellX = std::max(0, nMinCellX);
    if (nMinCellX >= FRAME_GRID_COLS) return vIndices;
    int nMaxCellX = std::ceil((x - mnMinX + r) * mfGridElementWidthInv);
    nMaxCellX = std::min(FRAME_GRID_COLS - 1, nMaxCellX);
    if (nMaxCellX < 0) return vIndices;
    int nMinCellY = std::floor((y - mnMinY - r) * mfGridElementHeightInv);
    nMinCellY = std::max(0, nMinCellY);
    if (nMinCellY >= FRAME_GRID_ROWS) return vIndices;
    int nMaxCellY = std::ceil((y - mnMinY + r) * mfGridElementHeightInv);
    nMaxCellY = std::min(FRAME_GRID_ROWS - 1, nMaxCellY);
    if (nMaxCellY < 0) return vIndices;

    bool bCheckLevels = true, bSameLevel = false;
    if (minLevel == -1 && maxLevel == -1) bCheckLevels = false;
    else if (minLevel == maxLevel) bSameLevel = true;

    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
            const std::vector<size_t>& vCell = mGrid[ix][iy];
            for (size_t j = 0; j < vCell.size(); j++) {
                const KeyPoint& kpUn = mvKeysUn[vCell[j]];
                if (bCheckLevels && !bSameLevel) {
                    if (kpUn.octave < minLevel || kpUn.octave > maxLevel) continue;
                } else if (bSameLevel) {
                    if (kpUn.octave != minLevel) continue;
                }
                if (std::abs(kpUn.x - x) > r || std::abs(kpUn.y - y) > r) continue;
                vIndices.push_back(vCell[j]);
            }
        }
    return vIndices;
}

// src/ORBmatcher.cc:2384-2400 (bit-hack popcount, restated verbatim in spirit)
int ORBmatcher::DescriptorDistance(const uint8_t* a, const uint8_t* b) {
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        uint32_t pa, pb;
        memcpy(&pa, a + 4 * i, 4);
        memcpy(&pb, b + 4 * i, 4);
        unsigned int v = pa ^ pb;
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

// src/ORBmatcher.cc:696-702  (float compared against the double literal 0.998)
float ORBmatcher::RadiusByViewingCos(float viewCos) {
    if (viewCos > 0.998) return 2.5;
    else return 4.0;
}

// include/ORBmatcher.h:71-145 == loop body of src/ORBmatcher.cc:384-465
int ORBmatcher::SearchByProjection_OnePoint(Frame& F, const MapPointView& mp, int mpId, float th) {
    const bool bFactor = th != 1.0;
    if (!mp.mbTrackInView) return -1;
    if (mp.isBad) return -1;
    const int nPredictedLevel = mp.mnTrackScaleLevel;
    float r = RadiusByViewingCos(mp.mTrackViewCos);
    if (bFactor) r *= th;
    std::vector<size_t> vNearIndices = F.GetFeaturesInArea(mp.mTrackProjX, mp.mTrackProjY,
                                                           r * F.mvScaleFactors[nPredictedLevel],
                                                           nPredictedLevel - 1, nPredictedLevel);
    if (vNearIndices.empty()) return -1;
    int bestDist = INT_MAX, bestLevel = -1, bestDist2 = INT_MAX, bestLevel2 = -1, bestIdx = -1;
    for (size_t idx : vNearIndices) {
        if (F.mvpMapPoints[idx] >= 0) continue;
        const int dist = DescriptorDistance(mp.desc, &F.mDescriptors[idx * 32]);
        if (dist < bestDist) {
            bestDist2 = bestDist; bestDist = dist;
            bestLevel2 = bestLevel; bestLevel = F.mvKeysUn[idx].octave;
            bestIdx = idx;
        } else if (dist < bestDist2) {
            bestLevel2 = F.mvKeysUn[idx].octave;
            bestDist2 = dist;
        }
    }
    if (bestDist <= TH_HIGH) {
        if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) return -1;
        F.mvpMapPoints[bestIdx] = mpId;
        F.mvpMatchScore[bestIdx] = bestDist;
        return bestIdx;
    }
    return -1;
}

int ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPointView>& vpMapPoints, float th) {
    int nmatches = 0;
    for (size_t iMP = 0; iMP < vpMapPoints.size(); iMP++)
        if (SearchByProjection_OnePoint(F, vpMapPoints[iMP], (int)iMP, th) >= 0) nmatches++;
    return nmatches;
}

// src/ORBmatcher.cc:2081-2202
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float* lastWorldPos, float th) {
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    const float* T = CurrentFrame.mTcw;
    for (size_t i = 0; i < LastFrame.mvpMapPoints.size(); i++) {
        if (LastFrame.mvpMapPoints[i] < 0) continue;
        if (LastFrame.mvbOutlier[i]) continue;
        const float* X = lastWorldPos + 3 * i;
        // cv::Mat x3Dc = Rcw*x3Dw+tcw: one cv::gemm(Rcw, x3Dw, 1, tcw, 1), small-matrix path (cvGemmRxPlusT, oracle.hpp)
        float xc3[3];
        cvGemmRxPlusT(T, 4, X, T + 3, 4, xc3);
        const float xc = xc3[0], yc = xc3[1];
        const float invzc = 1.0 / xc3[2];
        float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
        float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
        if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
        if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
        int nPredictedOctave = LastFrame.mvKeysUn[i].octave;
        float radius = th * CurrentFrame.mvScaleFactors[nPredictedOctave];
        std::vector<size_t> vIndices2 = CurrentFrame.GetFeaturesInArea(u, v, radius, nPredictedOctave - 1, nPredictedOctave + 1);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = &LastFrame.mDescriptors[i * 32];
        int bestDist = INT_MAX, bestIdx2 = -1;
        for (size_t i2 : vIndices2) {
            if (CurrentFrame.mvpMapPoints[i2] >= 0) continue;
            int dist = DescriptorDistance(dMP, &CurrentFrame.mDescriptors[i2 * 32]);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {
            CurrentFrame.mvpMapPoints[bestIdx2] = LastFrame.mvpMapPoints[i];
            CurrentFrame.mvpMatchScore[bestIdx2] = bestDist;
            nmatches++;
            if (mbCheckOrientation) {
                float rot = LastFrame.mvKeysUn[i].angle - CurrentFrame.mvKeysUn[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = std::round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) {
                    CurrentFrame.mvpMapPoints[idx] = -1;
                    CurrentFrame.mvpMatchScore[idx] = 999;
                    nmatches--;
                }
    }
    return nmatches;
}

// src/ORBmatcher.cc:51-140: SearchByProjection(F, MPs, th) with the window of a good feature (goodAtFrameId == F.mnId)
// scaled by th_good and of the rest by th_rest; F.mvpMatchScore is NOT written by this variant.
int ORBmatcher::SearchByProjection_GoodFeature(Frame& F, const std::vector<MapPointView>& vpMapPoints, const uint8_t* good,
                                               float th_good, float th_rest) {
    int nmatches = 0;
    const bool bFactor = th_rest != 1.0;
    for (size_t iMP = 0; iMP < vpMapPoints.size(); iMP++) {
        const MapPointView& mp = vpMapPoints[iMP];
        if (!mp.mbTrackInView) continue;
        if (mp.isBad) continue;
        const int nPredictedLevel = mp.mnTrackScaleLevel;
        float r = RadiusByViewingCos(mp.mTrackViewCos);
        if (good[iMP]) r *= th_good;
        else if (bFactor) r *= th_rest;
        std::vector<size_t> vNearIndices = F.GetFeaturesInArea(mp.mTrackProjX, mp.mTrackProjY, r * F.mvScaleFactors[nPredictedLevel],
                                                               nPredictedLevel - 1, nPredictedLevel);
        if (vNearIndices.empty()) continue;
        int bestDist = INT_MAX, bestLevel = -1, bestDist2 = INT_MAX, bestLevel2 = -1, bestIdx = -1;
        for (size_t idx : vNearIndices) {
            if (F.mvpMapPoints[idx] >= 0) continue;
            const int dist = DescriptorDistance(mp.desc, &F.mDescriptors[idx * 32]);
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist;
                bestLevel2 = bestLevel; bestLevel = F.mvKeysUn[idx].octave;
                bestIdx = idx;
            } else if (dist < bestDist2) {
                bestLevel2 = F.mvKeysUn[idx].octave;
                bestDist2 = dist;
            }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;
            F.mvpMapPoints[bestIdx] = (int)iMP;
            nmatches++;
        }
    }
    return nmatches;
}

// src/ORBmatcher.cc:143-274: SearchByProjection(Cur, Last, th) with radius th_good / th_rest per last-frame map point;
// no match score is stored or reset.
int ORBmatcher::SearchByProjection_GoodFeature(Frame& CurrentFrame, const Frame& LastFrame, const float* lastWorldPos,
                                               const uint8_t* good, float th_good, float th_rest) {
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    const float* T = CurrentFrame.mTcw;
    for (size_t i = 0; i < LastFrame.mvpMapPoints.size(); i++) {
        if (LastFrame.mvpMapPoints[i] < 0) continue;
        if (LastFrame.mvbOutlier[i]) continue;
        const float* X = lastWorldPos + 3 * i;
        float xc3[3];
        cvGemmRxPlusT(T, 4, X, T + 3, 4, xc3);
        const float xc = xc3[0], yc = xc3[1];
        const float invzc = 1.0 / xc3[2];
        float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
        float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
        if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
        if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
        int nPredictedOctave = LastFrame.mvKeysUn[i].octave;
        float radius;
        if (good[i]) radius = th_good * CurrentFrame.mvScaleFactors[nPredictedOctave];
        else radius = th_rest * CurrentFrame.mvScaleFactors[nPredictedOctave];
        std::vector<size_t> vIndices2 = CurrentFrame.GetFeaturesInArea(u, v, radius, nPredictedOctave - 1, nPredictedOctave + 1);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = &LastFrame.mDescriptors[i * 32];
        int bestDist = INT_MAX, bestIdx2 = -1;
        for (size_t i2 : vIndices2) {
            if (CurrentFrame.mvpMapPoints[i2] >= 0) continue;
            int dist = DescriptorDistance(dMP, &CurrentFrame.mDescriptors[i2 * 32]);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {
            CurrentFrame.mvpMapPoints[bestIdx2] = LastFrame.mvpMapPoints[i];
            nmatches++;
            if (mbCheckOrientation) {
                float rot = LastFrame.mvKeysUn[i].angle - CurrentFrame.mvKeysUn[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = std::round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) {
                    CurrentFrame.mvpMapPoints[idx] = -1;
                    nmatches--;
                }
    }
    return nmatches;
}

// src/ORBmatcher.cc:276-379: SearchByProjection(F, MPs, th) under a wall-clock budget; a non-positive budget matches
// nothing, and with an unlimited budget (the only mode the GPU path offers) the loop is the plain one.
int ORBmatcher::SearchByProjection_Budget(Frame& F, const std::vector<MapPointView>& vpMapPoints, float th, double time_constr) {
    if (time_constr <= 0) return 0;
    return SearchByProjection(F, vpMapPoints, th);
}

// src/ORBmatcher.cc:1172-1287
int ORBmatcher::SearchForInitialization(const Frame& F1, const Frame& F2, std::vector<float>& vbPrevMatched,
                                        std::vector<int>& vnMatches12, int windowSize) {
    int nmatches = 0;
    vnMatches12.assign(F1.mvKeysUn.size(), -1);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<int> vMatchedDistance(F2.mvKeysUn.size(), INT_MAX);
    std::vector<int> vnMatches21(F2.mvKeysUn.size(), -1);
    for (size_t i1 = 0; i1 < F1.mvKeysUn.size(); i1++) {
        const KeyPoint& kp1 = F1.mvKeysUn[i1];
        int level1 = kp1.octave;
        if (level1 > 0) continue;
        std::vector<size_t> vIndices2 = F2.GetFeaturesInArea(vbPrevMatched[2 * i1], vbPrevMatched[2 * i1 + 1],
                                                             windowSize, level1, level1);
        if (vIndices2.empty()) continue;
        const uint8_t* d1 = &F1.mDescriptors[i1 * 32];
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (size_t i2 : vIndices2) {
            int dist = DescriptorDistance(d1, &F2.mDescriptors[i2 * 32]);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * mfNNratio) {
                if (vnMatches21[bestIdx2] >= 0) {
                    vnMatches12[vnMatches21[bestIdx2]] = -1;
                    nmatches--;
                }
                vnMatches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (mbCheckOrientation) {
                    float rot = F1.mvKeysUn[i1].angle - F2.mvKeysUn[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(i1);
                }
            }
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; nmatches--; }
        }
    }
    for (size_t i1 = 0; i1 < vnMatches12.size(); i1++)
        if (vnMatches12[i1] >= 0) {
            vbPrevMatched[2 * i1] = F2.mvKeysUn[vnMatches12[i1]].x;
            vbPrevMatched[2 * i1 + 1] = F2.mvKeysUn[vnMatches12[i1]].y;
        }
    return nmatches;
}

// src/ORBmatcher.cc:979-1086.  F1.mvpMapPoints[i1] >= 0 plays the role of a non-NULL, non-bad MapPoint.
int ORBmatcher::WindowSearch(const Frame& F1, const Frame& F2, int windowSize, std::vector<int>& vpMapPointMatches2,
                             int minScaleLevel, int maxScaleLevel) {
    int nmatches = 0;
    vpMapPointMatches2.assign(F2.mvpMapPoints.size(), -1);
    std::vector<int> vnMatches21(F2.mvKeysUn.size(), -1);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    const bool bMinLevel = minScaleLevel > 0;
    const bool bMaxLevel = maxScaleLevel < INT_MAX;
    for (size_t i1 = 0; i1 < F1.mvpMapPoints.size(); i1++) {
        if (F1.mvpMapPoints[i1] < 0) continue;
        const KeyPoint& kp1 = F1.mvKeysUn[i1];
        int level1 = kp1.octave;
        if (bMinLevel && level1 < minScaleLevel) continue;
        if (bMaxLevel && level1 > maxScaleLevel) continue;
        std::vector<size_t> vIndices2 = F2.GetFeaturesInArea(kp1.x, kp1.y, windowSize, level1, level1);
        if (vIndices2.empty()) continue;
        const uint8_t* d1 = &F1.mDescriptors[i1 * 32];
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (size_t i2 : vIndices2) {
            if (vpMapPointMatches2[i2] >= 0) continue;
            int dist = DescriptorDistance(d1, &F2.mDescriptors[i2 * 32]);
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= bestDist2 * mfNNratio && bestDist <= TH_HIGH) {
            vpMapPointMatches2[bestIdx2] = F1.mvpMapPoints[i1];
            vnMatches21[bestIdx2] = i1;
            nmatches++;
            float rot = F1.mvKeysUn[i1].angle - F2.mvKeysUn[bestIdx2].angle;
            if (rot < 0.0) rot += 360.0f;
            int bin = std::round(rot * factor);
            if (bin == HISTO_LENGTH) bin = 0;
            rotHist[bin].push_back(bestIdx2);
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) {
                    vpMapPointMatches2[idx] = -1;
                    vnMatches21[idx] = -1;
                    nmatches--;
                }
    }
    return nmatches;
}

// src/ORBmatcher.cc:1089-1168: SearchByProjection(Frame& F1, Frame& F2, int windowSize, vector<MapPoint*>& vpMapPointMatches2)
// (TrackPreviousFrame, src/Tracking.cc:1373-1376).  f1World = MapPoint::GetWorldPos of F1.mvpMapPoints[i1]; ids >= 0 are
// non-NULL, non-bad map points.  F2.mvpMapPoints is only read; F2.mvpMatchScore is written.
int ORBmatcher::SearchByProjection(const Frame& F1, Frame& F2, const float* f1World, int windowSize, std::vector<int>& vpMapPointMatches2) {
    vpMapPointMatches2 = F2.mvpMapPoints;
    std::set<int> spMapPointsAlreadyFound(vpMapPointMatches2.begin(), vpMapPointMatches2.end());
    int nmatches = 0;
    const float* T = F2.mTcw;
    for (size_t i1 = 0; i1 < F1.mvpMapPoints.size(); i1++) {
        const int pMP1 = F1.mvpMapPoints[i1];
        if (pMP1 < 0) continue;
        if (spMapPointsAlreadyFound.count(pMP1)) continue;
        const int level1 = F1.mvKeysUn[i1].octave;
        const float* X = f1World + 3 * i1;
        float xc3[3];                                    // Rc2w*x3Dw+tc2w: see SearchByProjection(Cur, Last) above
        cvGemmRxPlusT(T, 4, X, T + 3, 4, xc3);
        const float xc2 = xc3[0], yc2 = xc3[1];
        const float invzc2 = 1.0 / xc3[2];
        float u2 = F2.fx * xc2 * invzc2 + F2.cx;
        float v2 = F2.fy * yc2 * invzc2 + F2.cy;
        std::vector<size_t> vIndices2 = F2.GetFeaturesInArea(u2, v2, windowSize, level1, level1);
        if (vIndices2.empty()) continue;
        const uint8_t* d1 = &F1.mDescriptors[i1 * 32];
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (size_t i2 : vIndices2) {
            if (vpMapPointMatches2[i2] >= 0) continue;
            int dist = DescriptorDistance(d1, &F2.mDescriptors[i2 * 32]);
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (static_cast<float>(bestDist) <= static_cast<float>(bestDist2) * mfNNratio && bestDist <= TH_HIGH) {
            vpMapPointMatches2[bestIdx2] = pMP1;
            F2.mvpMatchScore[bestIdx2] = bestDist;
            nmatches++;
        }
    }
    return nmatches;
}

// src/ORBmatcher.cc:2338-2379
void ORBmatcher::ComputeThreeMaxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = histo[i].size();
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) ind3 = -1;
}

// src/ORBmatcher.cc:724-853: SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches).  The DBoW2
// FeatureVector (std::map<NodeId, vector<unsigned>>, filled feature by feature in index order by
// Thirdparty/DBoW2/DBoW2/FeatureVector.cpp:addFeature) is rebuilt from the node id of every feature; a negative node id
// leaves the feature out.  KF.mvpMapPoints = pKF->GetMapPointMatches() with bad points as -1.
int ORBmatcher::SearchByBoW(const Frame& KF, const int* kfNode, const Frame& F, const int* fNode, std::vector<int>& vpMapPointMatches) {
    const std::vector<int>& vpMapPointsKF = KF.mvpMapPoints;
    vpMapPointMatches.assign(F.mvpMapPoints.size(), -1);
    std::map<int, std::vector<unsigned>> vFeatVecKF, vFeatVecF;
    for (int i = 0; i < KF.N; i++) if (kfNode[i] >= 0) vFeatVecKF[kfNode[i]].push_back((unsigned)i);
    for (int i = 0; i < F.N; i++) if (fNode[i] >= 0) vFeatVecF[fNode[i]].push_back((unsigned)i);
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    auto KFit = vFeatVecKF.begin(), KFend = vFeatVecKF.end();
    auto Fit = vFeatVecF.begin(), Fend = vFeatVecF.end();
    while (KFit != KFend && Fit != Fend) {
        if (KFit->first == Fit->first) {
            const std::vector<unsigned>& vIndicesKF = KFit->second;
            const std::vector<unsigned>& vIndicesF = Fit->second;
            for (size_t iKF = 0; iKF < vIndicesKF.size(); iKF++) {
                const unsigned realIdxKF = vIndicesKF[iKF];
                const int pMP = vpMapPointsKF[realIdxKF];
                if (pMP < 0) continue;                                   // NULL or isBad()
                const uint8_t* dKF = &KF.mDescriptors[(size_t)realIdxKF * 32];
                int bestDist1 = INT_MAX, bestIdxF = -1, bestDist2 = INT_MAX;
                for (size_t iF = 0; iF < vIndicesF.size(); iF++) {
                    const unsigned realIdxF = vIndicesF[iF];
                    if (vpMapPointMatches[realIdxF] >= 0) continue;
                    const int dist = DescriptorDistance(dKF, &F.mDescriptors[(size_t)realIdxF * 32]);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = (int)realIdxF; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (bestDist1 <= TH_LOW) {
                    if ((float)bestDist1 < mfNNratio * (float)bestDist2) {
                        vpMapPointMatches[bestIdxF] = pMP;
                        if (mbCheckOrientation) {
                            float rot = KF.mvKeysUn[realIdxKF].angle - F.mvKeysUn[bestIdxF].angle;
                            if (rot < 0.0) rot += 360.0f;
                            int bin = std::round(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back(bestIdxF);
                        }
                        nmatches++;
                    }
                }
            }
            KFit++;
            Fit++;
        } else if (KFit->first < Fit->first) {
            KFit = vFeatVecKF.lower_bound(Fit->first);
        } else {
            Fit = vFeatVecF.lower_bound(KFit->first);
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { vpMapPointMatches[idx] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// src/ORBmatcher.cc:2204-2336: SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound,
// float th, int ORBdist) (Tracking::Relocalization).  KF.mvpMapPoints = pKF->GetMapPointMatches(), kfSkip[i] != 0 stands
// for pMP->isBad() || sAlreadyFound.count(pMP); kfWorld / kfMinDist / kfMpDesc = GetWorldPos / GetMinDistanceInvariance /
// GetDescriptor of that map point; KF.mvKeysUn[i].angle = pKF->GetKeyPointUn(i).angle.
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& KF, const uint8_t* kfSkip, const float* kfWorld,
                                   const float* kfMinDist, const uint8_t* kfMpDesc, float th, int ORBdist) {
    int nmatches = 0;
    const float* T = CurrentFrame.mTcw;
    // const cv::Mat Ow = -Rcw.t()*tcw (:2210): cvNegRtTimesT (oracle.hpp)
    float Ow[3];
    cvNegRtTimesT(T, 4, T + 3, 4, Ow);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    for (size_t i = 0; i < KF.mvpMapPoints.size(); i++) {
        const int pMP = KF.mvpMapPoints[i];
        if (pMP < 0) continue;
        if (kfSkip && kfSkip[i]) continue;
        const float* X = kfWorld + 3 * i;
        float xc3[3];                                    // Rcw*x3Dw+tcw: see SearchByProjection(Cur, Last) above
        cvGemmRxPlusT(T, 4, X, T + 3, 4, xc3);
        const float xc = xc3[0], yc = xc3[1];
        const float invzc = 1.0 / xc3[2];
        float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
        float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
        if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
        if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
        // predicted scale level: cv::norm(x3Dw - Ow) (double accumulation, as in isInFrustum below)
        const float minDistance = kfMinDist[i];
        const float PO[3] = {X[0] - Ow[0], X[1] - Ow[1], X[2] - Ow[2]};
        float dist3D = (float)std::sqrt((double)PO[0] * PO[0] + (double)PO[1] * PO[1] + (double)PO[2] * PO[2]);
        float ratio = dist3D / minDistance;
        auto it = std::lower_bound(CurrentFrame.mvScaleFactors.begin(), CurrentFrame.mvScaleFactors.end(), ratio);
        const int nPredictedLevel = std::min((int)(it - CurrentFrame.mvScaleFactors.begin()), (int)CurrentFrame.mvScaleFactors.size() - 1);
        float radius = th * CurrentFrame.mvScaleFactors[nPredictedLevel];
        std::vector<size_t> vIndices2 = CurrentFrame.GetFeaturesInArea(u, v, radius, nPredictedLevel - 1, nPredictedLevel + 1);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = kfMpDesc + 32 * i;
        int bestDist = INT_MAX, bestIdx2 = -1;
        for (size_t i2 : vIndices2) {
            if (CurrentFrame.mvpMapPoints[i2] >= 0) continue;
            int dist = DescriptorDistance(dMP, &CurrentFrame.mDescriptors[i2 * 32]);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= ORBdist) {
            CurrentFrame.mvpMapPoints[bestIdx2] = pMP;
            CurrentFrame.mvpMatchScore[bestIdx2] = bestDist;
            nmatches++;
            if (mbCheckOrientation) {
                float rot = KF.mvKeysUn[i].angle - CurrentFrame.mvKeysUn[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = std::round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) {
                    CurrentFrame.mvpMapPoints[idx] = -1;
                    CurrentFrame.mvpMatchScore[idx] = 999;
                    nmatches--;
                }
    }
    return nmatches;
}

// src/Frame.cc:166-227.  cv::Mat arithmetic restated: Rcw*P+tcw is one small-matrix cv::gemm (cvGemmRxPlusT, oracle.hpp);
// cv::norm and Mat::dot of CV_32F vectors accumulate double products and return double.
bool isInFrustum(const float* Rcw, const float* tcw, const float* Ow, float fx, float fy, float cx, float cy, int mnMinX,
                 int mnMaxX, int mnMinY, int mnMaxY, const std::vector<float>& mvScaleFactors, const float* P,
                 const float* Pn, float minDistance, float maxDistance, float viewingCosLimit, float& u, float& v,
                 int& level, float& viewCos) {
    float Pc[3];
    cvGemmRxPlusT(Rcw, 3, P, tcw, 1, Pc);           // const cv::Mat Pc = mRcw*P+mtcw (:174)
    const float PcX = Pc[0], PcY = Pc[1], PcZ = Pc[2];
    if (PcZ < 0.0) return false;
    const float invz = 1.0 / PcZ;
    u = fx * PcX * invz + cx;
    v = fy * PcY * invz + cy;
    if (u < mnMinX || u > mnMaxX) return false;
    if (v < mnMinY || v > mnMaxY) return false;
    float PO[3];
    for (int r = 0; r < 3; r++) PO[r] = P[r] - Ow[r];
    double n2 = 0;
    for (int r = 0; r < 3; r++) n2 += (double)PO[r] * PO[r];
    const float dist = std::sqrt(n2);
    if (dist < minDistance || dist > maxDistance) return false;
    double dot = 0;
    for (int r = 0; r < 3; r++) dot += (double)PO[r] * Pn[r];
    viewCos = dot / dist;
    if (viewCos < viewingCosLimit) return false;
    const float ratio = dist / minDistance;
    int nPredictedLevel = (int)(std::lower_bound(mvScaleFactors.begin(), mvScaleFactors.end(), ratio) - mvScaleFactors.begin());
    if (nPredictedLevel >= (int)mvScaleFactors.size()) nPredictedLevel = (int)mvScaleFactors.size() - 1;
    level = nPredictedLevel;
    return true;
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:196-262) on the N observation descriptors of one map point:
// returns BestIdx (-1 when there is no observation: the reference returns early and keeps mDescriptor).
int distinctiveDescriptor(const uint8_t* descriptors, int N) {
    if (N <= 0) return -1;
    std::vector<float> Distances((size_t)N * N);
    for (int i = 0; i < N; i++) {
        Distances[(size_t)i * N + i] = 0;
        for (int j = i + 1; j < N; j++) {
            int distij = ORBmatcher::DescriptorDistance(descriptors + (size_t)32 * i, descriptors + (size_t)32 * j);
            Distances[(size_t)i * N + j] = distij;
            Distances[(size_t)j * N + i] = distij;
        }
    }
    int BestMedian = INT_MAX;
    int BestIdx = 0;
    for (int i = 0; i < N; i++) {
        std::vector<int> vDists(Distances.begin() + (size_t)i * N, Distances.begin() + (size_t)(i + 1) * N);
        std::sort(vDists.begin(), vDists.end());
        int median = vDists[0.5 * (N - 1)];
        if (median < BestMedian) {
            BestMedian = median;
            BestIdx = i;
        }
    }
    return BestIdx;
}

// Frame::UndistortKeyPoints (src/Frame.cc:389-423): cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) on the
// N x 2 CV_32F keypoint positions.  OpenCV is not vendored; cvUndistortPoints (imgproc/src/undistort.cpp, 2.4.13 and
// 3.4.1 agree for a k1,k2,p1,p2[,k3] model) restated: everything in double, K and the coefficients widened from
// CV_32F, 5 fixed-point iterations, R = I, P = K (so RR = K and ww = 1), result narrowed to float.  The caller handles
// the reference's early return (mDistCoef[0] == 0 -> mvKeysUn = mvKeys).
void undistortPoints(const float* src, int n, float fxf, float fyf, float cxf, float cyf, const float* dist, int ndist,
                     float* dst) {
    double k[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < ndist && i < 8; i++) k[i] = dist[i];
    const double fx = fxf, fy = fyf, cx = cxf, cy = cyf;
    const double ifx = 1. / fx, ify = 1. / fy;
    for (int i = 0; i < n; i++) {
        double x = src[2 * i], y = src[2 * i + 1];
        x = (x - cx) * ifx;
        y = (y - cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x);
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        const double xx = fx * x + 0. * y + cx;
        const double yy = 0. * x + fy * y + cy;
        const double ww = 1. / (0. * x + 0. * y + 1.);
        dst[2 * i] = (float)(xx * ww);
        dst[2 * i + 1] = (float)(yy * ww);
    }
}

}  // namespace oracle
