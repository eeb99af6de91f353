// oracle/orb_extract_oracle.hpp — TEST INFRASTRUCTURE (CPU oracle), never linked into the product.
//
// CPU restatement of ORBextractor (reference: ORB-SLAM2/src/ORBextractor.cpp, cited per function).
// Parity status: the reference ships no golden vectors; this restatement is pinned
//   (1) against OpenCV 4.13 (cv2) for the third-party primitives (tests/test_oracle_cv2.py), and
//   (2) against the reference's own ORBextractor.cpp compiled UNMODIFIED from /root/reference over a
//       mini-OpenCV shim (oracle/_ref, tests/test_oracle_vs_ref.py; fixtures in tests/golden/).
// Deliberate, documented deviation: the octree size-tie at ORBextractor.cpp:690 is broken by node
// creation order (canonical) instead of heap address; oracle/_ref reproduces the same order by
// running the reference under a monotonic bump allocator (address order == creation order).
#pragma once
#include "cv_prims.hpp"
#include <list>
#include <utility>

namespace orbo {

struct KeyPoint {            // binary layout of cv::KeyPoint (28 bytes)
    float x, y, size, angle, response;
    int octave, class_id;
};

static const int kPattern[1024] = {
#include "orb_pattern.inc"
};

static const int PATCH_SIZE = 31, HALF_PATCH_SIZE = 15, EDGE_THRESHOLD = 19;

struct Image {
    int w = 0, h = 0;
    std::vector<uint8_t> px;
    uint8_t* row(int y) { return px.data() + (size_t)y * w; }
    const uint8_t* row(int y) const { return px.data() + (size_t)y * w; }
    void alloc(int W, int H) { w = W; h = H; px.assign((size_t)W * H, 0); }
};

struct RawKp { float x, y; float response; };   // coordinates relative to minBorder (as in :829-834)

struct LevelDebug {
    std::vector<RawKp> toDistribute;   // cell-loop output in reference order (input of the octree)
    std::vector<KeyPoint> kps;         // after octree + border offset + angle (level coordinates)
};

struct Extractor {
    int nfeatures; double scaleFactor; int nlevels, iniThFAST, minThFAST;
    int tieMode = 0;          // 0 canonical (creation order), 1 faithful-to-this-process heap address
    bool legacyBlur = false;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
    std::vector<int> mnFeaturesPerLevel, umax;
    std::vector<Image> pyramid, blurred;
    std::vector<LevelDebug> dbg;

    // ORBextractor::ORBextractor  (ORBextractor.cpp:412-472)
    Extractor(int nf, float sf, int nl, int ini, int mn)
        : nfeatures(nf), scaleFactor(sf), nlevels(nl), iniThFAST(ini), minThFAST(mn)
    {
        mvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels);
        mvScaleFactor[0] = 1.0f; mvLevelSigma2[0] = 1.0f;
        for (int i = 1; i < nlevels; i++) {
            mvScaleFactor[i] = (float)(mvScaleFactor[i - 1] * scaleFactor);   // float*double -> float (:423)
            mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
        }
        mvInvScaleFactor.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
        for (int i = 0; i < nlevels; i++) {
            mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
            mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
        }
        mnFeaturesPerLevel.resize(nlevels);
        float factor = (float)(1.0f / scaleFactor);                                   // :438
        float nDesired = (float)(nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels)));
        int sum = 0;
        for (int level = 0; level < nlevels - 1; level++) {
            mnFeaturesPerLevel[level] = cvp::cv_round(nDesired);
            sum += mnFeaturesPerLevel[level];
            nDesired *= factor;
        }
        mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sum, 0);

        // circular patch row ends (:456-471)
        umax.resize(HALF_PATCH_SIZE + 1);
        int v, v0, vmax = cvp::cv_floor(HALF_PATCH_SIZE * sqrt(2.f) / 2 + 1);
        int vmin = cvp::cv_ceil(HALF_PATCH_SIZE * sqrt(2.f) / 2);
        const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
        for (v = 0; v <= vmax; ++v) umax[v] = cvp::cv_round(sqrt(hp2 - v * v));
        for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
            while (umax[v0] == umax[v0 + 1]) ++v0;
            umax[v] = v0;
            ++v0;
        }
        pyramid.resize(nlevels); blurred.resize(nlevels); dbg.resize(nlevels);
    }

    // ORBextractor::ComputePyramid (:1123-1148).  The 19-px copyMakeBorder is never read on the
    // hot path, so levels are stored border-less.
    void ComputePyramid(const uint8_t* img, int w, int h, size_t stride)
    {
        for (int level = 0; level < nlevels; ++level) {
            float scale = mvInvScaleFactor[level];
            int sw = cvp::cv_round((float)w * scale), sh = cvp::cv_round((float)h * scale);
            pyramid[level].alloc(sw, sh);
            if (level != 0)
                cvp::resize_linear_u8(pyramid[level - 1].px.data(), pyramid[level - 1].w, pyramid[level - 1].h,
                                      pyramid[level - 1].w, pyramid[level].px.data(), sw, sh, sw);
            else
                for (int y = 0; y < h; y++) memcpy(pyramid[0].row(y), img + (size_t)y * stride, w);
        }
    }

    // ---- octree (ORBextractor.cpp:483-770) ------------------------------------------------
    struct Node {
        int ULx, ULy, URx, URy, BLx, BLy, BRx, BRy;
        std::vector<RawKp> vKeys;
        std::list<Node>::iterator lit;
        bool bNoMore = false;
        long seq = 0;
    };
    long seqCounter = 0;

    // ExtractorNode::DivideNode (:483-539)
    static void DivideNode(const Node& p, Node& n1, Node& n2, Node& n3, Node& n4)
    {
        const int halfX = (int)ceil(static_cast<float>(p.URx - p.ULx) / 2);
        const int halfY = (int)ceil(static_cast<float>(p.BRy - p.ULy) / 2);
        n1.ULx = p.ULx; n1.ULy = p.ULy; n1.URx = p.ULx + halfX; n1.URy = p.ULy;
        n1.BLx = p.ULx; n1.BLy = p.ULy + halfY; n1.BRx = p.ULx + halfX; n1.BRy = p.ULy + halfY;
        n2.ULx = n1.URx; n2.ULy = n1.URy; n2.URx = p.URx; n2.URy = p.URy;
        n2.BLx = n1.BRx; n2.BLy = n1.BRy; n2.BRx = p.URx; n2.BRy = p.ULy + halfY;
        n3.ULx = n1.BLx; n3.ULy = n1.BLy; n3.URx = n1.BRx; n3.URy = n1.BRy;
        n3.BLx = p.BLx; n3.BLy = p.BLy; n3.BRx = n1.BRx; n3.BRy = p.BLy;
        n4.ULx = n3.URx; n4.ULy = n3.URy; n4.URx = n2.BRx; n4.URy = n2.BRy;
        n4.BLx = n3.BRx; n4.BLy = n3.BRy; n4.BRx = p.BRx; n4.BRy = p.BRy;
        for (size_t i = 0; i < p.vKeys.size(); i++) {
            const RawKp& kp = p.vKeys[i];
            if (kp.x < n1.URx) {
                if (kp.y < n1.BRy) n1.vKeys.push_back(kp); else n3.vKeys.push_back(kp);
            } else if (kp.y < n1.BRy) n2.vKeys.push_back(kp);
            else n4.vKeys.push_back(kp);
        }
        if (n1.vKeys.size() == 1) n1.bNoMore = true;
        if (n2.vKeys.size() == 1) n2.bNoMore = true;
        if (n3.vKeys.size() == 1) n3.bNoMore = true;
        if (n4.vKeys.size() == 1) n4.bNoMore = true;
    }

    struct SizeNode { int size; Node* node; };

    void sortSizeNodes(std::vector<SizeNode>& v) const
    {
        if (tieMode == 0)
            std::sort(v.begin(), v.end(), [](const SizeNode& a, const SizeNode& b) {
                return a.size != b.size ? a.size < b.size : a.node->seq < b.node->seq; });
        else
            std::sort(v.begin(), v.end(), [](const SizeNode& a, const SizeNode& b) {
                return a.size != b.size ? a.size < b.size : a.node < b.node; });
    }

    // ORBextractor::DistributeOctTree (:541-770)
    std::vector<RawKp> DistributeOctTree(const std::vector<RawKp>& keys, int minX, int maxX, int minY, int maxY, int N)
    {
        int nIni = (int)round(static_cast<float>(maxX - minX) / (maxY - minY));
        if (keys.empty()) return std::vector<RawKp>();
        if (nIni < 1) nIni = 1;   // aspect < 0.5: the reference indexes an empty vector (crash); defined as one root.
        const float hX = static_cast<float>(maxX - minX) / nIni;
        std::list<Node> lNodes;
        std::vector<Node*> ini(nIni);
        for (int i = 0; i < nIni; i++) {
            Node ni;
            ni.ULx = (int)(hX * static_cast<float>(i)); ni.ULy = 0;
            ni.URx = (int)(hX * static_cast<float>(i + 1)); ni.URy = 0;
            ni.BLx = ni.ULx; ni.BLy = maxY - minY;
            ni.BRx = ni.URx; ni.BRy = maxY - minY;
            ni.seq = seqCounter++;
            lNodes.push_back(ni);
            ini[i] = &lNodes.back();
        }
        for (size_t i = 0; i < keys.size(); i++) ini[(int)(keys[i].x / hX)]->vKeys.push_back(keys[i]);

        auto lit = lNodes.begin();
        while (lit != lNodes.end()) {
            if (lit->vKeys.size() == 1) { lit->bNoMore = true; lit++; }
            else if (lit->vKeys.empty()) lit = lNodes.erase(lit);
            else lit++;
        }

        bool bFinish = false;
        std::vector<SizeNode> vSize;
        auto addChild = [&](Node& c, int* nToExpand) {
            if (c.vKeys.size() > 0) {
                c.seq = seqCounter++;
                lNodes.push_front(c);
                if (c.vKeys.size() > 1) {
                    if (nToExpand) (*nToExpand)++;
                    vSize.push_back(SizeNode{(int)c.vKeys.size(), &lNodes.front()});
                    lNodes.front().lit = lNodes.begin();
                }
            }
        };
        while (!bFinish) {
            int prevSize = (int)lNodes.size();
            lit = lNodes.begin();
            int nToExpand = 0;
            vSize.clear();
            while (lit != lNodes.end()) {
                if (lit->bNoMore) { lit++; continue; }
                Node n1, n2, n3, n4;
                DivideNode(*lit, n1, n2, n3, n4);
                addChild(n1, &nToExpand); addChild(n2, &nToExpand);
                addChild(n3, &nToExpand); addChild(n4, &nToExpand);
                lit = lNodes.erase(lit);
            }
            if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) {
                bFinish = true;
            } else if (((int)lNodes.size() + nToExpand * 3) > N) {
                while (!bFinish) {
                    prevSize = (int)lNodes.size();
                    std::vector<SizeNode> vPrev = vSize;
                    vSize.clear();
                    sortSizeNodes(vPrev);
                    for (int j = (int)vPrev.size() - 1; j >= 0; j--) {
                        Node n1, n2, n3, n4;
                        DivideNode(*vPrev[j].node, n1, n2, n3, n4);
                        addChild(n1, nullptr); addChild(n2, nullptr);
                        addChild(n3, nullptr); addChild(n4, nullptr);
                        lNodes.erase(vPrev[j].node->lit);
                        if ((int)lNodes.size() >= N) break;
                    }
                    if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) bFinish = true;
                }
            }
        }
        std::vector<RawKp> result;
        result.reserve(nfeatures);
        for (auto it = lNodes.begin(); it != lNodes.end(); it++) {
            const std::vector<RawKp>& nk = it->vKeys;
            const RawKp* best = &nk[0];
            float maxResponse = best->response;
            for (size_t k = 1; k < nk.size(); k++)
                if (nk[k].response > maxResponse) { best = &nk[k]; maxResponse = nk[k].response; }
            result.push_back(*best);
        }
        return result;
    }

    // IC_Angle (:79-106)
    float IC_Angle(const Image& im, float px, float py) const
    {
        int m_01 = 0, m_10 = 0;
        const int cx = cvp::cv_round(px), cy = cvp::cv_round(py);
        const uint8_t* center = im.row(cy) + cx;
        for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
        const int step = im.w;
        for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
            int v_sum = 0, d = umax[v];
            for (int u = -d; u <= d; ++u) {
                int val_plus = center[u + v * step], val_minus = center[u - v * step];
                v_sum += (val_plus - val_minus);
                m_10 += u * (val_plus + val_minus);
            }
            m_01 += v * v_sum;
        }
        return cvp::fast_atan2((float)m_01, (float)m_10);
    }

    // ComputeKeyPointsOctTree (:772-863)
    void ComputeKeyPointsOctTree(std::vector<std::vector<KeyPoint>>& all)
    {
        all.resize(nlevels);
        const float W = 30;
        for (int level = 0; level < nlevels; ++level) {
            const Image& im = pyramid[level];
            const int minBorderX = EDGE_THRESHOLD - 3, minBorderY = minBorderX;
            const int maxBorderX = im.w - EDGE_THRESHOLD + 3, maxBorderY = im.h - EDGE_THRESHOLD + 3;
            std::vector<RawKp>& toDist = dbg[level].toDistribute;
            toDist.clear();
            const float width = (float)(maxBorderX - minBorderX), height = (float)(maxBorderY - minBorderY);
            const int nCols = (int)(width / W), nRows = (int)(height / W);
            if (nCols < 1 || nRows < 1) {   // level smaller than one 30-px cell: the reference divides by zero (UB);
                all[level].clear();         // defined here (and in the CUDA path) as "no keypoints on this level".
                continue;
            }
            const int wCell = (int)ceil(width / nCols), hCell = (int)ceil(height / nRows);
            std::vector<cvp::FastKp> cell;
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
                    const int x0 = (int)iniX, y0 = (int)iniY, cw = (int)maxX - x0, ch = (int)maxY - y0;
                    cvp::fast9_16_nms(im.row(y0) + x0, cw, ch, im.w, iniThFAST, cell);
                    if (cell.empty()) cvp::fast9_16_nms(im.row(y0) + x0, cw, ch, im.w, minThFAST, cell);
                    for (const auto& k : cell)
                        toDist.push_back(RawKp{(float)k.x + j * wCell, (float)k.y + i * hCell, (float)k.score});
                }
            }
            std::vector<RawKp> sel = DistributeOctTree(toDist, minBorderX, maxBorderX, minBorderY, maxBorderY,
                                                       mnFeaturesPerLevel[level]);
            const int scaledPatchSize = (int)(PATCH_SIZE * mvScaleFactor[level]);
            std::vector<KeyPoint>& kps = all[level];
            kps.clear();
            for (const auto& r : sel) {
                KeyPoint k;
                k.x = r.x + minBorderX; k.y = r.y + minBorderY; k.size = (float)scaledPatchSize;
                k.angle = -1.f; k.response = r.response; k.octave = level; k.class_id = -1;
                kps.push_back(k);
            }
        }
        for (int level = 0; level < nlevels; ++level)
            for (auto& k : all[level]) k.angle = IC_Angle(pyramid[level], k.x, k.y);
    }

    // computeOrbDescriptor (:110-149) — no FMA contraction (file is compiled -ffp-contract=off).
    static void computeOrbDescriptor(const KeyPoint& kpt, const Image& img, uint8_t* desc)
    {
        const float factorPI = (float)(3.14159265358979323846 / 180.f);
        float angle = (float)kpt.angle * factorPI;
        float a = (float)cos(angle), b = (float)sin(angle);   // `using namespace std` => float overloads
        const int cx = cvp::cv_round(kpt.x), cy = cvp::cv_round(kpt.y);
        const uint8_t* center = img.row(cy) + cx;
        const int step = img.w;
        const int* pattern = kPattern;
        auto get = [&](int idx) -> int {
            const int px = pattern[2 * idx], py = pattern[2 * idx + 1];
            return center[cvp::cv_round(px * b + py * a) * step + cvp::cv_round(px * a - py * b)];
        };
        for (int i = 0; i < 32; ++i, pattern += 32) {
            int val = 0;
            for (int bit = 0; bit < 8; bit++) {
                int t0 = get(2 * bit), t1 = get(2 * bit + 1);
                val |= (t0 < t1) << bit;
            }
            desc[i] = (uint8_t)val;
        }
    }

    // ORBextractor::operator() (:1053-1117)
    void run(const uint8_t* img, int w, int h, size_t stride, std::vector<KeyPoint>& out, std::vector<uint8_t>& desc)
    {
        out.clear(); desc.clear();
        if (!img || w <= 0 || h <= 0) return;
        ComputePyramid(img, w, h, stride);
        std::vector<std::vector<KeyPoint>> all;
        ComputeKeyPointsOctTree(all);
        int n = 0;
        for (int level = 0; level < nlevels; ++level) n += (int)all[level].size();
        desc.resize((size_t)n * 32);
        int offset = 0;
        for (int level = 0; level < nlevels; ++level) {
            std::vector<KeyPoint>& kps = all[level];
            dbg[level].kps = kps;
            // the reference blurs only levels that have keypoints (:1092-1098); the oracle blurs all
            // levels so that stage-parity tests can compare every level.
            blurred[level].alloc(pyramid[level].w, pyramid[level].h);
            cvp::gaussian_blur7_u8(pyramid[level].px.data(), pyramid[level].w, pyramid[level].h, pyramid[level].w,
                                   blurred[level].px.data(), pyramid[level].w, legacyBlur);
            if (kps.empty()) continue;
            for (size_t i = 0; i < kps.size(); i++)
                computeOrbDescriptor(kps[i], blurred[level], desc.data() + (size_t)(offset + i) * 32);
            offset += (int)kps.size();
            if (level != 0) {
                float scale = mvScaleFactor[level];
                for (auto& k : kps) { k.x *= scale; k.y *= scale; }
            }
            out.insert(out.end(), kps.begin(), kps.end());
        }
    }
};

}  // namespace orbo
