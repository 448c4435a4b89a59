// TEST INFRASTRUCTURE (see oracle.hpp).  The non-default keypoint distribution mode of SURVEY 8(f)-4: the quadtree of
// ORB-SLAM2 (Mur-Artal & Tardos, "ORB-SLAM2", IEEE T-RO 2017; ORBextractor::ComputeKeyPointsOctTree / DistributeOctTree /
// ExtractorNode::DivideNode of the published ORB_SLAM2 sources).  /root/reference is the ORB-SLAM1-derived GF fork and holds
// NO quadtree code, so there is no reference file:line to cite and nothing to pin against: **parity unpinned** for this mode,
// the CUDA path is checked against this restatement only.  Two things the published code leaves open are fixed here:
//   * std::sort of (key count, node pointer) pairs breaks ties by heap address; here ties go by node creation order
//     (one of the executions the original can produce);
//   * a level may return up to N + 3 keypoints and the frame therefore more than nfeatures; rows beyond nfeatures are
//     dropped at the end (the C ABI's output arrays hold nfeatures rows).
#include "oracle.hpp"
#include <algorithm>
#include <cmath>
#include <list>
#include <utility>
#include <vector>

namespace oracle {

namespace {

struct ExtractorNode {
    int ULx = 0, ULy = 0, URx = 0, BLy = 0;        // UL, UR = (URx, ULy), BL = (ULx, BLy), BR = (URx, BLy)
    std::vector<KeyPoint> vKeys;
    bool bNoMore = false;
    int id = 0;                                    // creation order (tie-break of the size sort)
    std::list<ExtractorNode>::iterator lit;

    void DivideNode(ExtractorNode& n1, ExtractorNode& n2, ExtractorNode& n3, ExtractorNode& n4) const {
        const int halfX = std::ceil((float)(URx - ULx) / 2);
        const int halfY = std::ceil((float)(BLy - ULy) / 2);
        n1.ULx = ULx; n1.ULy = ULy; n1.URx = ULx + halfX; n1.BLy = ULy + halfY;
        n2.ULx = n1.URx; n2.ULy = ULy; n2.URx = URx; n2.BLy = ULy + halfY;
        n3.ULx = ULx; n3.ULy = n1.BLy; n3.URx = n1.URx; n3.BLy = BLy;
        n4.ULx = n1.URx; n4.ULy = n1.BLy; n4.URx = URx; n4.BLy = BLy;
        for (const KeyPoint& kp : vKeys) {
            if (kp.x < n1.URx) { if (kp.y < n1.BLy) n1.vKeys.push_back(kp); else n3.vKeys.push_back(kp); }
            else if (kp.y < n1.BLy) n2.vKeys.push_back(kp);
            else n4.vKeys.push_back(kp);
        }
        if (n1.vKeys.size() == 1) n1.bNoMore = true;
        if (n2.vKeys.size() == 1) n2.bNoMore = true;
        if (n3.vKeys.size() == 1) n3.bNoMore = true;
        if (n4.vKeys.size() == 1) n4.bNoMore = true;
    }
};

std::vector<KeyPoint> DistributeOctTree(const std::vector<KeyPoint>& vToDistributeKeys, int minX, int maxX, int minY, int maxY, int N) {
    const int nIni = std::round((float)(maxX - minX) / (maxY - minY));
    std::vector<KeyPoint> vResultKeys;
    if (nIni <= 0) return vResultKeys;             // portrait levels narrower than half their height: the original divides by zero
    const float hX = (float)(maxX - minX) / nIni;
    std::list<ExtractorNode> lNodes;
    std::vector<ExtractorNode*> vpIniNodes(nIni);
    int nextId = 0;
    for (int i = 0; i < nIni; i++) {
        ExtractorNode ni;
        ni.ULx = (int)(hX * (float)i); ni.URx = (int)(hX * (float)(i + 1)); ni.ULy = 0; ni.BLy = maxY - minY;
        ni.id = nextId++;
        lNodes.push_back(ni);
        vpIniNodes[i] = &lNodes.back();
    }
    for (const KeyPoint& kp : vToDistributeKeys) {
        const int slot = (int)(kp.x / hX);
        if (slot < 0 || slot >= nIni) continue;    // cannot happen for keypoints inside [0, maxX - minX)
        vpIniNodes[slot]->vKeys.push_back(kp);
    }
    for (auto lit = lNodes.begin(); lit != lNodes.end();) {
        if (lit->vKeys.size() == 1) { lit->bNoMore = true; ++lit; }
        else if (lit->vKeys.empty()) lit = lNodes.erase(lit);
        else ++lit;
    }
    bool bFinish = false;
    std::vector<std::pair<std::pair<int, int>, ExtractorNode*>> vSizeAndPointerToNode;      // ((size, id), node)
    auto addChild = [&](ExtractorNode& n, int* nToExpand) {
        if (n.vKeys.size() > 0) {
            n.id = nextId++;
            lNodes.push_front(n);
            if (n.vKeys.size() > 1) {
                if (nToExpand) (*nToExpand)++;
                vSizeAndPointerToNode.push_back({{(int)n.vKeys.size(), lNodes.front().id}, &lNodes.front()});
                lNodes.front().lit = lNodes.begin();
            }
        }
    };
    while (!bFinish) {
        int prevSize = (int)lNodes.size();
        auto lit = lNodes.begin();
        int nToExpand = 0;
        vSizeAndPointerToNode.clear();
        while (lit != lNodes.end()) {
            if (lit->bNoMore) { ++lit; continue; }
            ExtractorNode n1, n2, n3, n4;
            lit->DivideNode(n1, n2, n3, n4);
            addChild(n1, &nToExpand); addChild(n2, &nToExpand); addChild(n3, &nToExpand); addChild(n4, &nToExpand);
            lit = lNodes.erase(lit);
        }
        if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) {
            bFinish = true;
        } else if ((int)lNodes.size() + nToExpand * 3 > N) {
            while (!bFinish) {
                prevSize = (int)lNodes.size();
                auto vPrev = vSizeAndPointerToNode;
                vSizeAndPointerToNode.clear();
                std::sort(vPrev.begin(), vPrev.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
                for (int j = (int)vPrev.size() - 1; j >= 0; j--) {
                    ExtractorNode n1, n2, n3, n4;
                    vPrev[j].second->DivideNode(n1, n2, n3, n4);
                    addChild(n1, nullptr); addChild(n2, nullptr); addChild(n3, nullptr); addChild(n4, nullptr);
                    lNodes.erase(vPrev[j].second->lit);
                    if ((int)lNodes.size() >= N) break;
                }
                if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) bFinish = true;
            }
        }
    }
    vResultKeys.reserve(lNodes.size());
    for (auto& node : lNodes) {                    // the best response of every node, first one on ties
        const KeyPoint* pKP = &node.vKeys[0];
        float maxResponse = pKP->response;
        for (size_t k = 1; k < node.vKeys.size(); k++)
            if (node.vKeys[k].response > maxResponse) { pKP = &node.vKeys[k]; maxResponse = node.vKeys[k].response; }
        vResultKeys.push_back(*pKP);
    }
    return vResultKeys;
}

}  // namespace

// ORB-SLAM2 ORBextractor::ComputeKeyPointsOctTree on this extractor's pyramid and features-per-level table
int ORBextractor::ComputeKeyPointsOctTree(std::vector<std::vector<KeyPoint>>& allKeypoints) {
    allKeypoints.assign(nlevels, {});
    const float W = 30;
    for (int level = 0; level < nlevels; ++level) {
        Level& L = mvImagePyramid[level];
        const int minBorderX = 16 - 3, minBorderY = minBorderX;
        const int maxBorderX = L.w - 16 + 3, maxBorderY = L.h - 16 + 3;
        std::vector<KeyPoint> vToDistributeKeys;
        const float width = (float)(maxBorderX - minBorderX), height = (float)(maxBorderY - minBorderY);
        const int nCols = (int)(width / W), nRows = (int)(height / W);
        if (nCols <= 0 || nRows <= 0) return -2;   // the original divides by zero
        const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
        for (int i = 0; i < nRows; i++) {
            const float iniY = (float)(minBorderY + i * hCell);
            float maxY = iniY + hCell + 6;
            if (iniY >= maxBorderY - 3) continue;
            if (maxY > maxBorderY) maxY = (float)maxBorderY;
            for (int j = 0; j < nCols; j++) {
                const float iniX = (float)(minBorderX + j * wCell);
                float maxX = iniX + wCell + 6;
                if (iniX >= maxBorderX - 6) continue;
                if (maxX > maxBorderX) maxX = (float)maxBorderX;
                const int x0 = (int)iniX, x1 = (int)maxX, y0 = (int)iniY, y1 = (int)maxY;
                const uint8_t* cell = L.roi() + (ptrdiff_t)y0 * L.step + x0;
                std::vector<KeyPoint> vKeysCell;
                FAST9_16(cell, x1 - x0, y1 - y0, L.step, fastTh, true, vKeysCell);
                if (vKeysCell.empty()) FAST9_16(cell, x1 - x0, y1 - y0, L.step, minFastTh, true, vKeysCell);
                for (KeyPoint& kp : vKeysCell) {
                    kp.x += j * wCell; kp.y += i * hCell;
                    vToDistributeKeys.push_back(kp);
                }
            }
        }
        std::vector<KeyPoint>& keypoints = allKeypoints[level];
        keypoints = DistributeOctTree(vToDistributeKeys, minBorderX, maxBorderX, minBorderY, maxBorderY, mnFeaturesPerLevel[level]);
        const int scaledPatchSize = 31 * mvScaleFactor[level];
        for (KeyPoint& kp : keypoints) {
            kp.x += minBorderX; kp.y += minBorderY;
            kp.octave = level;
            kp.size = scaledPatchSize;
        }
    }
    // rows beyond nfeatures are dropped (see the header of this file)
    int total = 0;
    for (int level = 0; level < nlevels; ++level) {
        const int room = std::max(nfeatures - total, 0);
        if ((int)allKeypoints[level].size() > room) allKeypoints[level].resize(room);
        total += (int)allKeypoints[level].size();
    }
    for (int level = 0; level < nlevels; ++level) {
        Level& L = mvImagePyramid[level];
        for (auto& kp : allKeypoints[level]) kp.angle = ic_angle_at(L.roi(), L.step, kp.x, kp.y);
    }
    return 0;
}

}  // namespace oracle
