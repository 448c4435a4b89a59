// Non-default keypoint distribution (what north_star names "quadtree keypoint distribution"): the quadtree of
// ORB-SLAM2 (ORBextractor::ComputeKeyPointsOctTree / DistributeOctTree / ExtractorNode::DivideNode of the published ORB_SLAM2
// sources).  The GF-ORB-SLAM reference has no such code (it distributes by the cell quota, k_quota / k_cell_retain); the
// CUDA path is checked against oracle/orb_quadtree.cpp, which fixes the two points the published code leaves open (size-sort
// ties by node creation order; rows beyond nfeatures dropped).  Selected by gforb_extractor_set_distribution(h, 1).
//
// Detection reuses k_fast_nms / k_cell_emit on the 30-pixel cell grid of this mode (cell interiors tile [16, w-16) x [16, h-16)
// exactly like the quota grid; the FAST fallback is "cell empty at iniThFAST -> minThFAST").  k_quadtree then runs one warp
// per (frame, level): the level's keys are gathered in the published order (cells row-major, pixels row-major), a node owns
// a contiguous range of that array, DivideNode is a stable 4-way partition of the range done by the 32 lanes (ballot ranks),
// and the node list (push_front / erase / front-to-back walk) is a doubly linked list of slots allocated in creation order,
// so a slot index is the creation id that breaks ties of the size sort.  Control flow is uniform over the warp; lane 0
// writes the node fields.
#pragma once

namespace gforb {

struct QtLevel {                 // per level, planned on the host (float arithmetic of the published code)
    int nIni;                    // round((maxX - minX) / (maxY - minY)) initial nodes
    float hX;                    // (maxX - minX) / nIni
    int rootW, rootH;            // maxX - minX, maxY - minY  (minX = minY = 13, max = size - 13)
    int iniUL[5];                // (int)(hX * i), i = 0 .. nIni
    int N;                       // mnFeaturesPerLevel
    int keyBase, keyCap;         // slice of the per-frame key arrays (entries)
    int nodeBase, nodeCap;       // slice of the per-frame node array
};

struct QtNode {
    short ULx, ULy, URx, BLy;
    int begin, count;            // range of the level's key array
    short prev, next;            // list links (-1 = none)
    short noMore, pad;
};

static const int QT_MIN_BORDER = 13;         // EDGE_THRESHOLD - 3

// Stable 4-way partition of keys[begin, begin + n) by class (0..3) into tmp and back; returns the four class counts in cnt[].
// cls(key) must be uniform code; all 32 lanes call.
template <class F>
__device__ __forceinline__ void qt_partition4(uint32_t* __restrict__ keys, uint32_t* __restrict__ tmp, int begin, int n, F cls,
                                              int (&cnt)[4], int lane) {
    const unsigned full = 0xffffffffu, lt = (1u << lane) - 1u;
    int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    for (int i0 = 0; i0 < n; i0 += 32) {
        const int i = i0 + lane;
        const int c = i < n ? cls(keys[begin + i]) : -1;
        c0 += __popc(__ballot_sync(full, c == 0)); c1 += __popc(__ballot_sync(full, c == 1));
        c2 += __popc(__ballot_sync(full, c == 2)); c3 += __popc(__ballot_sync(full, c == 3));
    }
    cnt[0] = c0; cnt[1] = c1; cnt[2] = c2; cnt[3] = c3;
    int b0 = 0, b1 = c0, b2 = c0 + c1, b3 = c0 + c1 + c2;
    for (int i0 = 0; i0 < n; i0 += 32) {
        const int i = i0 + lane;
        const uint32_t k = i < n ? keys[begin + i] : 0u;
        const int c = i < n ? cls(k) : -1;
        const unsigned m0 = __ballot_sync(full, c == 0), m1 = __ballot_sync(full, c == 1);
        const unsigned m2 = __ballot_sync(full, c == 2), m3 = __ballot_sync(full, c == 3);
        if (c == 0) tmp[begin + b0 + __popc(m0 & lt)] = k;
        else if (c == 1) tmp[begin + b1 + __popc(m1 & lt)] = k;
        else if (c == 2) tmp[begin + b2 + __popc(m2 & lt)] = k;
        else if (c == 3) tmp[begin + b3 + __popc(m3 & lt)] = k;
        b0 += __popc(m0); b1 += __popc(m1); b2 += __popc(m2); b3 += __popc(m3);
    }
    __syncwarp();
    for (int i = lane; i < n; i += 32) keys[begin + i] = tmp[begin + i];
    __syncwarp();
}

// key = score << 24 | qy << 12 | qx with (qx, qy) relative to the level's (minBorderX, minBorderY) = (13, 13)
__device__ __forceinline__ int qt_x(uint32_t k) { return (int)(k & 0xfff); }
__device__ __forceinline__ int qt_y(uint32_t k) { return (int)((k >> 12) & 0xfff); }

struct QtList {                  // the std::list<ExtractorNode> of DistributeOctTree
    QtNode* nodes;
    int head, tail, size, nAlloc, cap;
};

// DivideNode of node p + the caller's bookkeeping (push the non-empty children to the front in the order n1..n4, record the
// children with more than one key in vsz, erase p).  Returns false when the node array is full.
__device__ __forceinline__ bool qt_divide(QtList& Lst, int p, uint32_t* keys, uint32_t* tmp, int2* vsz, int& nVsz, int& nToExpand,
                                          int lane) {
    const QtNode P = Lst.nodes[p];
    const int midX = P.ULx + ((P.URx - P.ULx + 1) >> 1);          // ceil((float)(URx - ULx) / 2)
    const int midY = P.ULy + ((P.BLy - P.ULy + 1) >> 1);
    int cnt[4];
    qt_partition4(keys, tmp, P.begin, P.count,
                  [=](uint32_t k) { return (qt_x(k) < midX ? 0 : 1) + (qt_y(k) < midY ? 0 : 2); }, cnt, lane);
    int off = P.begin;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        if (cnt[c] > 0) {
            if (Lst.nAlloc >= Lst.cap) return false;
            const int id = Lst.nAlloc++;
            if (lane == 0) {
                QtNode n;
                n.ULx = (short)((c & 1) ? midX : P.ULx); n.URx = (short)((c & 1) ? P.URx : midX);
                n.ULy = (short)((c & 2) ? midY : P.ULy); n.BLy = (short)((c & 2) ? P.BLy : midY);
                n.begin = off; n.count = cnt[c];
                n.noMore = cnt[c] == 1; n.pad = 0;
                n.prev = -1; n.next = (short)Lst.head;
                Lst.nodes[id] = n;
                if (Lst.head >= 0) Lst.nodes[Lst.head].prev = (short)id;
                if (cnt[c] > 1) vsz[nVsz] = make_int2(cnt[c], id);
            }
            if (Lst.tail < 0) Lst.tail = id;
            Lst.head = id;
            Lst.size++;
            if (cnt[c] > 1) { nVsz++; nToExpand++; }
            __syncwarp();
        }
        off += cnt[c];
    }
    // lNodes.erase(p): p is never the head here when a child was pushed, but may be when every child was empty (cannot
    // happen: count >= 2) -- handled generally
    const int pp = Lst.nodes[p].prev, pn = P.next;
    __syncwarp();
    if (lane == 0) {
        if (pp >= 0) Lst.nodes[pp].next = (short)pn;
        if (pn >= 0) Lst.nodes[pn].prev = (short)pp;
    }
    if (pp < 0) Lst.head = pn;
    if (pn < 0) Lst.tail = pp;
    Lst.size--;
    __syncwarp();
    return true;
}

// One CTA (8 warps) per frame, warp w takes the levels w, w + 8, ...; after the levels of a frame are done the rows beyond
// nfeatures are dropped (levels in order).  lvlKp / lvlCount are what k_orient and k_describe read in either mode.
__global__ void __launch_bounds__(256) k_quadtree(const LevelPlan* __restrict__ levels, const QtLevel* __restrict__ qlv, int nLevels,
                                                  const CellPlan* __restrict__ cells, int nCellsTotal,
                                                  const uint32_t* __restrict__ cand, size_t frameCand,
                                                  const int* __restrict__ cnt7, const int* __restrict__ cnt20, int iniTh,
                                                  uint32_t* __restrict__ keysAll, uint32_t* __restrict__ tmpAll, size_t frameKeys,
                                                  QtNode* __restrict__ nodesAll, int2* __restrict__ vszAll, size_t frameNodes,
                                                  uint32_t* __restrict__ lvlKp, size_t frameKp, int* __restrict__ lvlCount,
                                                  int nfeatures, int* __restrict__ errflag) {
    const unsigned full = 0xffffffffu;
    const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    for (int l = warp; l < nLevels; l += 8) {
        const LevelPlan L = levels[l];
        const QtLevel Q = qlv[l];
        uint32_t* keys = keysAll + (size_t)b * frameKeys + Q.keyBase;
        uint32_t* tmp = tmpAll + (size_t)b * frameKeys + Q.keyBase;
        QtList Lst;
        Lst.nodes = nodesAll + (size_t)b * frameNodes + Q.nodeBase;
        Lst.cap = Q.nodeCap; Lst.head = Lst.tail = -1; Lst.size = 0; Lst.nAlloc = 0;
        int2* vsz = vszAll + ((size_t)b * frameNodes + Q.nodeBase) * 2;       // two lists of nodeCap entries
        int2* vprev = vsz + Q.nodeCap;
        // ---- vToDistributeKeys: cells row-major, the cell's FAST keys at iniTh if there are any, else at minTh
        int nKeys = 0;
        bool overflow = false;
        for (int c = 0; c < L.nCells; c++) {
            const int cellId = L.cellBase + c;
            const size_t ci = (size_t)b * nCellsTotal + cellId;
            const int n7 = cnt7[ci], n20 = cnt20[ci];
            if (n7 <= 0) continue;
            const uint32_t* list = cand + (size_t)b * frameCand + cells[cellId].candBase;
            const bool ini = n20 > 0;
            const int take = ini ? n20 : n7;
            if (nKeys + take > Q.keyCap) { overflow = true; break; }
            for (int i0 = 0; i0 < n7; i0 += 32) {
                const int i = i0 + lane;
                const uint32_t v = i < n7 ? list[i] : 0u;
                const bool keep = i < n7 && (!ini || (int)(v >> 24) >= iniTh);
                const unsigned m = __ballot_sync(full, keep);
                if (keep) keys[nKeys + __popc(m & lt)] = v - (uint32_t)(QT_MIN_BORDER | (QT_MIN_BORDER << 12));
                nKeys += __popc(m);
            }
        }
        __syncwarp();
        int nOut = 0;
        uint32_t* out = lvlKp + (size_t)b * frameKp + L.kpBase;
        if (!overflow && nKeys > 0) {
            // ---- initial nodes: stable partition by (int)(kp.x / hX)
            const float hX = Q.hX;
            const int nIni = Q.nIni;
            int cnt[4];
            qt_partition4(keys, tmp, 0, nKeys, [=](uint32_t k) { return min((int)__fdiv_rn((float)qt_x(k), hX), nIni - 1); }, cnt, lane);
            int off = 0;
            for (int i = 0; i < nIni; i++) {                 // push_back in order; empty ones are erased, single keys are final
                const int id = Lst.nAlloc++;
                if (cnt[i] > 0) {
                    if (lane == 0) {
                        QtNode n;
                        n.ULx = (short)Q.iniUL[i]; n.URx = (short)Q.iniUL[i + 1]; n.ULy = 0; n.BLy = (short)Q.rootH;
                        n.begin = off; n.count = cnt[i]; n.noMore = cnt[i] == 1; n.pad = 0;
                        n.prev = (short)Lst.tail; n.next = -1;
                        Lst.nodes[id] = n;
                        if (Lst.tail >= 0) Lst.nodes[Lst.tail].next = (short)id;
                    }
                    if (Lst.head < 0) Lst.head = id;
                    Lst.tail = id;
                    Lst.size++;
                    __syncwarp();
                }
                off += cnt[i];
            }
            const int N = Q.N;
            bool finish = false, full_ = false;
            int nVsz = 0;
            while (!finish) {
                const int prevSize = Lst.size;
                int nToExpand = 0;
                nVsz = 0;
                int it = Lst.head;
                while (it >= 0) {
                    const int nx = Lst.nodes[it].next;
                    if (!Lst.nodes[it].noMore) {
                        if (!qt_divide(Lst, it, keys, tmp, vsz, nVsz, nToExpand, lane)) { full_ = true; break; }
                    }
                    it = nx;
                }
                if (full_) break;
                if (Lst.size >= N || Lst.size == prevSize) {
                    finish = true;
                } else if (Lst.size + nToExpand * 3 > N) {
                    while (!finish) {
                        const int prev2 = Lst.size;
                        const int nPrev = nVsz;
                        // vPrev sorted ascending by (size, id) and walked from the back == descending order
                        for (int e = lane; e < nPrev; e += 32) {
                            const int2 me = vsz[e];
                            int rank = 0;
                            for (int o = 0; o < nPrev; o++) {
                                const int2 ot = vsz[o];
                                rank += (ot.x > me.x || (ot.x == me.x && ot.y > me.y)) ? 1 : 0;
                            }
                            vprev[rank] = me;
                        }
                        __syncwarp();
                        nVsz = 0;
                        int dummy = 0;
                        for (int j = 0; j < nPrev; j++) {
                            if (!qt_divide(Lst, vprev[j].y, keys, tmp, vsz, nVsz, dummy, lane)) { full_ = true; break; }
                            if (Lst.size >= N) break;
                        }
                        if (full_) break;
                        if (Lst.size >= N || Lst.size == prev2) finish = true;
                    }
                    if (full_) break;
                }
            }
            if (full_) overflow = true;
            // ---- the best response of every node (first one on ties), list order
            if (!overflow) {
                for (int it = Lst.head; it >= 0; it = Lst.nodes[it].next) {
                    const QtNode nd = Lst.nodes[it];
                    unsigned best = 0;                                  // score << 20 | (0xfffff - position)
                    for (int i = lane; i < nd.count; i += 32) {
                        const uint32_t k = keys[nd.begin + i];
                        best = max(best, ((k >> 24) << 20) | (0xfffffu - (unsigned)i));
                    }
                    best = __reduce_max_sync(full, best);
                    const uint32_t k = keys[nd.begin + (int)(0xfffffu - (best & 0xfffffu))];
                    if (lane == 0 && nOut < L.kpCap) out[nOut] = k + (uint32_t)(QT_MIN_BORDER | (QT_MIN_BORDER << 12));
                    nOut++;
                }
                if (nOut > L.kpCap) { nOut = L.kpCap; overflow = true; }
            }
        }
        if (overflow && lane == 0) atomicExch(errflag, GFORB_ERR_OVERFLOW);
        if (lane == 0) lvlCount[b * nLevels + l] = overflow ? 0 : nOut;
    }
    __syncthreads();
    if (threadIdx.x == 0) {                                  // rows beyond nfeatures are dropped, levels in order
        int total = 0;
        for (int l = 0; l < nLevels; l++) {
            const int room = max(nfeatures - total, 0);
            int n = lvlCount[b * nLevels + l];
            if (n > room) { n = room; lvlCount[b * nLevels + l] = n; }
            total += n;
        }
    }
}

}  // namespace gforb
