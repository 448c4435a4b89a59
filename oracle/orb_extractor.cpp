// TEST INFRASTRUCTURE ONLY (see oracle.hpp).  CPU restatement of ORB_SLAM::ORBextractor
// (src/ORBextractor.cc:131-201, 464-533, 535-767, 769-875, 922-998).
#include "oracle.hpp"
#include <algorithm>
#include <cmath>
#include <cstring>

namespace oracle {

static const int PATCH_SIZE = 31, HALF_PATCH_SIZE = 15, EDGE_THRESHOLD = 16;   // src/ORBextractor.cc:82-84
static const float HARRIS_K = 0.04f;

static const int bit_pattern_31_[256][4] = {
#include "../gf_orb_slam_b200/csrc/brief_pattern.inc"
};

// src/ORBextractor.cc:464-524
ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _scoreType, int _fastTh,
                           int _minFastTh, int _blurVariant, int _retainVariant)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), scoreType(_scoreType), fastTh(_fastTh),
      minFastTh(_minFastTh), blurVariant(_blurVariant), retainVariant(_retainVariant) {
    mvScaleFactor.resize(nlevels);
    mvScaleFactor[0] = 1;
    for (int i = 1; i < nlevels; i++) mvScaleFactor[i] = mvScaleFactor[i - 1] * scaleFactor;   // float*double->float

    float invScaleFactor = 1.0f / scaleFactor;
    mvInvScaleFactor.resize(nlevels);
    mvInvScaleFactor[0] = 1;
    for (int i = 1; i < nlevels; i++) mvInvScaleFactor[i] = mvInvScaleFactor[i - 1] * invScaleFactor;

    mvImagePyramid.resize(nlevels);
    mnFeaturesPerLevel.resize(nlevels);
    float factor = (float)(1.0 / scaleFactor);
    float nDesiredFeaturesPerScale = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sumFeatures = 0;
    for (int level = 0; level < nlevels - 1; level++) {
        mnFeaturesPerLevel[level] = cvRoundf(nDesiredFeaturesPerScale);
        sumFeatures += mnFeaturesPerLevel[level];
        nDesiredFeaturesPerScale *= factor;
    }
    mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sumFeatures, 0);

    umax.resize(HALF_PATCH_SIZE + 1);
    int v, v0, vmax = (int)std::floor(HALF_PATCH_SIZE * std::sqrt(2.f) / 2 + 1);
    int vmin = (int)std::ceil(HALF_PATCH_SIZE * std::sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (v = 0; v <= vmax; ++v) umax[v] = cvRound(std::sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
        while (umax[v0] == umax[v0 + 1]) ++v0;
        umax[v] = v0;
        ++v0;
    }
}

// src/ORBextractor.cc:922-998
void ORBextractor::ComputePyramid(const uint8_t* image, int w, int h, int step) {
    for (int level = 0; level < nlevels; ++level) {
        float scale = mvInvScaleFactor[level];
        int sw = cvRoundf((float)w * scale), sh = cvRoundf((float)h * scale);
        Level& L = mvImagePyramid[level];
        L.w = sw; L.h = sh; L.step = sw + EDGE_THRESHOLD * 2;
        L.buf.assign((size_t)L.step * (sh + EDGE_THRESHOLD * 2), 0);
        if (level != 0) {
            Level& P = mvImagePyramid[level - 1];
            // resize reads the *unblurred* previous level ROI, writes this level's ROI, then the ring is
            // filled with BORDER_REFLECT_101 (+ISOLATED) of this level
            std::vector<uint8_t> tmp((size_t)sw * sh);
            resizeLinear8u(P.roi(), P.w, P.h, P.step, tmp.data(), sw, sh, sw);
            copyMakeBorder101(tmp.data(), sw, sh, sw, L.buf.data(), L.step, EDGE_THRESHOLD);
        } else {
            copyMakeBorder101(image, w, h, step, L.buf.data(), L.step, EDGE_THRESHOLD);
        }
    }
}

// src/ORBextractor.cc:86-127
static void HarrisResponses(const uint8_t* img, int step, std::vector<KeyPoint>& pts, int blockSize, float harris_k) {
    int r = blockSize / 2;
    float scale = (1 << 2) * blockSize * 255.0f;
    scale = 1.0f / scale;
    float scale_sq_sq = scale * scale * scale * scale;
    for (auto& p : pts) {
        int x0 = cvRoundf(p.x - r), y0 = cvRoundf(p.y - r);
        const uint8_t* ptr0 = img + (ptrdiff_t)y0 * step + x0;
        int a = 0, b = 0, c = 0;
        for (int i = 0; i < blockSize; i++)
            for (int j = 0; j < blockSize; j++) {
                const uint8_t* ptr = ptr0 + i * step + j;
                int Ix = (ptr[1] - ptr[-1]) * 2 + (ptr[-step + 1] - ptr[-step - 1]) + (ptr[step + 1] - ptr[step - 1]);
                int Iy = (ptr[step] - ptr[-step]) * 2 + (ptr[step - 1] - ptr[-step - 1]) + (ptr[step + 1] - ptr[-step + 1]);
                a += Ix * Ix; b += Iy * Iy; c += Ix * Iy;
            }
        p.response = ((float)a * b - (float)c * c - harris_k * ((float)a + b) * ((float)a + b)) * scale_sq_sq;
    }
}

// src/ORBextractor.cc:131-158
static float IC_Angle(const uint8_t* roi, int step, float ptx, float pty, const std::vector<int>& u_max) {
    int m_01 = 0, m_10 = 0;
    const uint8_t* center = roi + (ptrdiff_t)cvRoundf(pty) * step + cvRoundf(ptx);
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
        int v_sum = 0, d = u_max[v];
        for (int u = -d; u <= d; ++u) {
            int val_plus = center[u + v * step], val_minus = center[u - v * step];
            v_sum += (val_plus - val_minus);
            m_10 += u * (val_plus + val_minus);
        }
        m_01 += v * v_sum;
    }
    return fastAtan2((float)m_01, (float)m_10);
}

// src/ORBextractor.cc:161-201
static const float factorPI = (float)(M_PI / 180.f);
static void computeOrbDescriptor(const KeyPoint& kpt, const uint8_t* roi, int step, uint8_t* desc) {
    float angle = (float)kpt.angle * factorPI;
    // `(float)cos(angle)` with a float argument under `using namespace std` (src/ORBextractor.cc:75) resolves to
    // std::cos(float) == libm cosf/sinf, NOT the double routines.
    float a = cosf(angle), b = sinf(angle);
    const uint8_t* center = roi + (ptrdiff_t)cvRoundf(kpt.y) * step + cvRoundf(kpt.x);
    auto get = [&](int x, int y) -> int {
        return center[cvRoundf(x * b + y * a) * step + cvRoundf(x * a - y * b)];
    };
    for (int i = 0; i < 32; ++i) {
        int val = 0;
        for (int k = 0; k < 8; k++) {
            const int* p = bit_pattern_31_[i * 8 + k];
            int t0 = get(p[0], p[1]), t1 = get(p[2], p[3]);
            val |= (t0 < t1) << k;
        }
        desc[i] = (uint8_t)val;
    }
}

// single-keypoint entry points for the known-answer tests (tests/test_oracle_kat.py): the orientation and the descriptor of
// a keypoint at (x, y) on a plane the caller supplies
float ic_angle_at(const uint8_t* img, int step, float x, float y) {
    static const std::vector<int> umax = ORBextractor(100, 1.2f, 1).umax;
    return IC_Angle(img, step, x, y, umax);
}
void describe_at(const uint8_t* img, int step, float x, float y, float angleDeg, uint8_t* desc) {
    KeyPoint kp{};
    kp.x = x; kp.y = y; kp.angle = angleDeg;
    computeOrbDescriptor(kp, img, step, desc);
}

// src/ORBextractor.cc:535-758.  Returns <0 where the reference would hit UB / a cv::Exception.
int ORBextractor::ComputeKeyPoints(std::vector<std::vector<KeyPoint>>& allKeypoints) {
    allKeypoints.assign(nlevels, {});
    float imageRatio = (float)mvImagePyramid[0].w / mvImagePyramid[0].h;
    for (int level = 0; level < nlevels; ++level) {
        Level& L = mvImagePyramid[level];
        const int nDesiredFeatures = mnFeaturesPerLevel[level];
        const int levelCols = std::sqrt((float)nDesiredFeatures / (5 * imageRatio));
        const int levelRows = imageRatio * levelCols;
        if (levelCols <= 0 || levelRows <= 0) return -2;   // reference: division by zero below
        const int minBorderX = EDGE_THRESHOLD, minBorderY = minBorderX;
        const int maxBorderX = L.w - EDGE_THRESHOLD, maxBorderY = L.h - EDGE_THRESHOLD;
        const int W = maxBorderX - minBorderX, H = maxBorderY - minBorderY;
        const int cellW = std::ceil((float)W / levelCols), cellH = std::ceil((float)H / levelRows);
        const int nCells = levelRows * levelCols;
        const int nfeaturesCell = std::ceil((float)nDesiredFeatures / nCells);

        std::vector<std::vector<std::vector<KeyPoint>>> cellKeyPoints(levelRows, std::vector<std::vector<KeyPoint>>(levelCols));
        std::vector<std::vector<int>> nToRetain(levelRows, std::vector<int>(levelCols, 0));
        std::vector<std::vector<int>> nTotal(levelRows, std::vector<int>(levelCols, 0));
        std::vector<std::vector<char>> bNoMore(levelRows, std::vector<char>(levelCols, 0));
        std::vector<int> iniXCol(levelCols), iniYRow(levelRows);
        int nNoMore = 0, nToDistribute = 0;
        float hY = cellH + 6;
        for (int i = 0; i < levelRows; i++) {
            const float iniY = minBorderY + i * cellH - 3;
            iniYRow[i] = iniY;
            if (i == levelRows - 1) {
                hY = maxBorderY + 3 - iniY;
                if (hY <= 0) continue;
            }
            float hX = cellW + 6;
            for (int j = 0; j < levelCols; j++) {
                float iniX;
                if (i == 0) { iniX = minBorderX + j * cellW - 3; iniXCol[j] = iniX; }
                else iniX = iniXCol[j];
                if (j == levelCols - 1) {
                    hX = maxBorderX + 3 - iniX;
                    if (hX <= 0) continue;
                }
                // Mat::rowRange/colRange on the level ROI: must stay inside [0,rows]x[0,cols] (CV_Assert otherwise)
                int x0 = (int)iniX, x1 = (int)(iniX + hX), y0 = (int)iniY, y1 = (int)(iniY + hY);
                if (x0 < 0 || y0 < 0 || x1 > L.w || y1 > L.h || x0 > x1 || y0 > y1) return -3;
                const uint8_t* cell = L.roi() + (ptrdiff_t)y0 * L.step + x0;
                std::vector<KeyPoint>& kc = cellKeyPoints[i][j];
                FAST9_16(cell, x1 - x0, y1 - y0, L.step, fastTh, true, kc);
                if (kc.size() <= 3) {
                    kc.clear();
                    FAST9_16(cell, x1 - x0, y1 - y0, L.step, minFastTh, true, kc);
                }
                if (scoreType == HARRIS_SCORE) HarrisResponses(cell, L.step, kc, 7, HARRIS_K);
                const int nKeys = kc.size();
                nTotal[i][j] = nKeys;
                if (nKeys > nfeaturesCell) { nToRetain[i][j] = nfeaturesCell; bNoMore[i][j] = 0; }
                else { nToRetain[i][j] = nKeys; nToDistribute += nfeaturesCell - nKeys; bNoMore[i][j] = 1; nNoMore++; }
            }
        }
        while (nToDistribute > 0 && nNoMore < nCells) {
            int nNewFeaturesCell = nfeaturesCell + std::ceil((float)nToDistribute / (nCells - nNoMore));
            nToDistribute = 0;
            for (int i = 0; i < levelRows; i++)
                for (int j = 0; j < levelCols; j++)
                    if (!bNoMore[i][j]) {
                        if (nTotal[i][j] > nNewFeaturesCell) { nToRetain[i][j] = nNewFeaturesCell; bNoMore[i][j] = 0; }
                        else {
                            nToRetain[i][j] = nTotal[i][j];
                            nToDistribute += nNewFeaturesCell - nTotal[i][j];
                            bNoMore[i][j] = 1; nNoMore++;
                        }
                    }
        }
        std::vector<KeyPoint>& keypoints = allKeypoints[level];
        keypoints.reserve(nDesiredFeatures * 2);
        const int scaledPatchSize = PATCH_SIZE * mvScaleFactor[level];
        for (int i = 0; i < levelRows; i++)
            for (int j = 0; j < levelCols; j++) {
                std::vector<KeyPoint>& keysCell = cellKeyPoints[i][j];
                retainBest(keysCell, nToRetain[i][j], retainVariant);
                if ((int)keysCell.size() > nToRetain[i][j]) keysCell.resize(nToRetain[i][j]);
                for (size_t k = 0; k < keysCell.size(); k++) {
                    keysCell[k].x += iniXCol[j];
                    keysCell[k].y += iniYRow[i];
                    keysCell[k].octave = level;
                    keysCell[k].size = scaledPatchSize;
                    keypoints.push_back(keysCell[k]);
                }
            }
        if ((int)keypoints.size() > nDesiredFeatures) {
            retainBest(keypoints, nDesiredFeatures, retainVariant);
            keypoints.resize(nDesiredFeatures);
        }
    }
    for (int level = 0; level < nlevels; ++level) {
        Level& L = mvImagePyramid[level];
        for (auto& kp : allKeypoints[level]) kp.angle = IC_Angle(L.roi(), L.step, kp.x, kp.y, umax);
    }
    return 0;
}

// src/ORBextractor.cc:769-875
int ORBextractor::operator()(const uint8_t* image, int w, int h, int step, std::vector<KeyPoint>& _keypoints,
                             std::vector<uint8_t>& _descriptors) {
    if (image == nullptr || w <= 0 || h <= 0) return 0;   // _image.empty() -> silent return
    ComputePyramid(image, w, h, step);
    std::vector<std::vector<KeyPoint>> allKeypoints;
    int rc = distribution == 1 ? ComputeKeyPointsOctTree(allKeypoints) : ComputeKeyPoints(allKeypoints);
    if (rc < 0) return rc;
    int nkeypoints = 0;
    for (int level = 0; level < nlevels; ++level) nkeypoints += (int)allKeypoints[level].size();
    _descriptors.assign((size_t)nkeypoints * 32, 0);
    _keypoints.clear();
    _keypoints.reserve(nkeypoints);
    lastAllKeypoints = allKeypoints;
    int offset = 0;
    for (int level = 0; level < nlevels; ++level) {
        std::vector<KeyPoint>& keypoints = allKeypoints[level];
        int nkeypointsLevel = (int)keypoints.size();
        if (nkeypointsLevel == 0) continue;
        Level& L = mvImagePyramid[level];
        gaussianBlur7x7InPlace(L.roi(), L.w, L.h, L.step, blurVariant);
        for (int i = 0; i < nkeypointsLevel; i++)
            computeOrbDescriptor(keypoints[i], L.roi(), L.step, _descriptors.data() + (size_t)(offset + i) * 32);
        offset += nkeypointsLevel;
        if (level != 0) {
            float scale = mvScaleFactor[level];
            for (auto& kp : keypoints) { kp.x *= scale; kp.y *= scale; }
        }
        _keypoints.insert(_keypoints.end(), keypoints.begin(), keypoints.end());
    }
    return 0;
}

}  // namespace oracle
