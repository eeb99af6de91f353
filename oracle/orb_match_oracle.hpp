// oracle/orb_match_oracle.hpp — TEST INFRASTRUCTURE (CPU oracle), never linked into the product.
//
// CPU restatement of the Hamming-matching half of the hot path.  Reference files followed:
//   ORB-SLAM2/src/ORBmatcher.cpp  :37-39 (constants) :63-155 :166-190 :205-348 :604-743 (SearchByBoW) :755-927 (SearchForTriangulation) :352-472 :962-1251 (Fuse) :1253-1500 (SearchBySim3) :1661-1789 :474-590 :1506-1659 :1792-1854
//   ORB-SLAM2/src/Frame.cpp       :241-257 :379-444 :540-748
//   ORB-SLAM2/src/MapPoint.cpp    :302-380 (ComputeDistinctiveDescriptors)
// PINNED: the reference ships no tests or golden vectors for these functions, so the reference itself was run here —
// ORBmatcher.cpp, Frame.cpp, KeyFrame.cpp, MapPoint.cpp, Map.cpp and the vendored DBoW2 compile UNMODIFIED over oracle/cvshim into
// oracle/_ref/liborb_refmatch.so (oracle/Makefile `refmatch`, harness oracle/refmatch_capi.cpp) — and its outputs on seeded scenes
// are committed as tests/golden/ref_match.npz (generator tests/golden/make_golden_match.py).  tests/test_ref_match_golden.py checks
// every function below against that fixture bit for bit (and the CUDA path against the same fixture under -m gpu).
// MapPoint*/Frame bookkeeping is flattened to plain arrays; the control flow follows the reference.
#pragma once
#include <cstdint>
#include <cmath>
#include <climits>
#include <vector>
#include <map>
#include <algorithm>
#include "orb_extract_oracle.hpp"

namespace orbo {

static const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;          // ORBmatcher.cpp:37-39
static const int FRAME_GRID_ROWS = 48, FRAME_GRID_COLS = 64;              // Frame.h:37-38

// ORBmatcher::DescriptorDistance (ORBmatcher.cpp:1838-1854): SWAR popcount over 8 x 32 bit.
static inline int DescriptorDistance(const uint8_t* a, const uint8_t* b)
{
    const int32_t* pa = (const int32_t*)a;
    const int32_t* pb = (const int32_t*)b;
    int dist = 0;
    for (int i = 0; i < 8; i++, pa++, pb++) {
        unsigned int v = *pa ^ *pb;
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

// MapPoint::ComputeDistinctiveDescriptors (MapPoint.cpp:302-380) on the N descriptor rows of one map point: the float
// distance matrix (:340-352), the per-row sort of the distances as ints and the median vDists[0.5*(N-1)] (:358-365),
// first smallest median wins (:368-372).  Returns the winning row, -1 when the point has no rows (:316-317, :330-331).
static inline int DistinctiveDescriptor(const uint8_t* rows, int N, int* median_out)
{
    if (N <= 0) { if (median_out) *median_out = -1; return -1; }
    std::vector<std::vector<float> > Distances(N, std::vector<float>(N, 0));
    for (int i = 0; i < N; i++) {
        Distances[i][i] = 0;
        for (int j = i + 1; j < N; j++) {
            int distij = DescriptorDistance(rows + (size_t)i * 32, rows + (size_t)j * 32);
            Distances[i][j] = distij;
            Distances[j][i] = distij;
        }
    }
    int BestMedian = INT_MAX, BestIdx = 0;
    for (int i = 0; i < N; i++) {
        std::vector<int> vDists(Distances[i].begin(), Distances[i].end());
        std::sort(vDists.begin(), vDists.end());
        int median = vDists[0.5 * (N - 1)];
        if (median < BestMedian) { BestMedian = median; BestIdx = i; }
    }
    if (median_out) *median_out = BestMedian;
    return BestIdx;
}

// Frame grid (Frame.cpp:241-257, :434-444; constructor lines :115-116 for the inverse cell sizes)
struct FrameGrid {
    int N = 0;
    const KeyPoint* kps = nullptr;        // undistorted keypoints (mvKeysUn)
    const uint8_t* desc = nullptr;
    float mnMinX, mnMinY, mnMaxX, mnMaxY, invW, invH;
    std::vector<int> cells[FRAME_GRID_COLS][FRAME_GRID_ROWS];

    void build(const KeyPoint* k, const uint8_t* d, int n, float minx, float miny, float maxx, float maxy)
    {
        kps = k; desc = d; N = n; mnMinX = minx; mnMinY = miny; mnMaxX = maxx; mnMaxY = maxy;
        invW = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(mnMaxX - mnMinX);
        invH = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(mnMaxY - mnMinY);
        for (auto& col : cells) for (auto& c : col) c.clear();
        for (int i = 0; i < N; i++) {
            int px = (int)round((kps[i].x - mnMinX) * invW);
            int py = (int)round((kps[i].y - mnMinY) * invH);
            if (px < 0 || px >= FRAME_GRID_COLS || py < 0 || py >= FRAME_GRID_ROWS) continue;
            cells[px][py].push_back(i);
        }
    }

    // Frame::GetFeaturesInArea (Frame.cpp:379-432)
    void GetFeaturesInArea(float x, float y, float r, int minLevel, int maxLevel, std::vector<int>& out) const
    {
        out.clear();
        const int nMinCellX = std::max(0, (int)floor((x - mnMinX - r) * invW));
        if (nMinCellX >= FRAME_GRID_COLS) return;
        const int nMaxCellX = std::min((int)FRAME_GRID_COLS - 1, (int)ceil((x - mnMinX + r) * invW));
        if (nMaxCellX < 0) return;
        const int nMinCellY = std::max(0, (int)floor((y - mnMinY - r) * invH));
        if (nMinCellY >= FRAME_GRID_ROWS) return;
        const int nMaxCellY = std::min((int)FRAME_GRID_ROWS - 1, (int)ceil((y - mnMinY + r) * invH));
        if (nMaxCellY < 0) return;
        const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                const std::vector<int>& vCell = cells[ix][iy];
                for (size_t j = 0; j < vCell.size(); j++) {
                    const KeyPoint& kp = kps[vCell[j]];
                    if (bCheckLevels) {
                        if (kp.octave < minLevel) continue;
                        if (maxLevel >= 0 && kp.octave > maxLevel) continue;
                    }
                    const float distx = kp.x - x, disty = kp.y - y;
                    if (fabs(distx) < r && fabs(disty) < r) out.push_back(vCell[j]);
                }
            }
    }
};

// ORBmatcher::ComputeThreeMaxima (ORBmatcher.cpp:1792-1833)
static inline void ComputeThreeMaxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3)
{
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = (int)histo[i].size();
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

// ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&) (ORBmatcher.cpp:205-348) on flattened arrays.
// FeatureVectors are std::map<NodeId, vector<unsigned>>; kf_has_mp[i] = "vpMapPointsKF[i] && !isBad()"; the result stores the
// KeyFrame feature index where the reference stores its MapPoint pointer.
static inline int SearchByBoW(const uint8_t* descKF, const float* angleKF, const uint8_t* kf_has_mp,
                              const std::map<unsigned, std::vector<unsigned> >& vFeatVecKF,
                              const uint8_t* descF, const float* angleF, int FN,
                              const std::map<unsigned, std::vector<unsigned> >& vFeatVecF,
                              float mfNNratio, bool mbCheckOrientation, std::vector<int>& vpMapPointMatches)
{
    vpMapPointMatches.assign(FN, -1);
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = HISTO_LENGTH / 360.0f;
    std::map<unsigned, std::vector<unsigned> >::const_iterator KFit = vFeatVecKF.begin(), Fit = vFeatVecF.begin();
    const std::map<unsigned, std::vector<unsigned> >::const_iterator KFend = vFeatVecKF.end(), Fend = vFeatVecF.end();
    while (KFit != KFend && Fit != Fend) {
        if (KFit->first == Fit->first) {
            const std::vector<unsigned> vIndicesKF = KFit->second;
            const std::vector<unsigned> vIndicesF = Fit->second;
            for (size_t iKF = 0; iKF < vIndicesKF.size(); iKF++) {
                const unsigned realIdxKF = vIndicesKF[iKF];
                if (!kf_has_mp[realIdxKF]) continue;
                const uint8_t* dKF = descKF + (size_t)realIdxKF * 32;
                int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
                for (size_t iF = 0; iF < vIndicesF.size(); iF++) {
                    const unsigned realIdxF = vIndicesF[iF];
                    if (vpMapPointMatches[realIdxF] >= 0) continue;
                    const int dist = DescriptorDistance(dKF, descF + (size_t)realIdxF * 32);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (bestDist1 <= TH_LOW) {
                    if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
                        vpMapPointMatches[bestIdxF] = (int)realIdxKF;
                        if (mbCheckOrientation) {
                            float rot = angleKF[realIdxKF] - angleF[bestIdxF];
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)round(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back(bestIdxF);
                        }
                        nmatches++;
                    }
                }
            }
            KFit++; Fit++;
        } else if (KFit->first < Fit->first) KFit = vFeatVecKF.lower_bound(Fit->first);
        else Fit = vFeatVecF.lower_bound(KFit->first);
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) { vpMapPointMatches[rotHist[i][j]] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12) (ORBmatcher.cpp:604-743) on flattened
// arrays: has_mp1/has_mp2 = "MapPoint exists and is not bad"; the result stores the KeyFrame-2 feature index.
static inline int SearchByBoW_KF(const uint8_t* Descriptors1, const float* angle1, const uint8_t* has_mp1, int n1,
                                 const std::map<unsigned, std::vector<unsigned> >& vFeatVec1,
                                 const uint8_t* Descriptors2, const float* angle2, const uint8_t* has_mp2, int n2,
                                 const std::map<unsigned, std::vector<unsigned> >& vFeatVec2,
                                 float mfNNratio, bool mbCheckOrientation, std::vector<int>& vpMatches12)
{
    vpMatches12.assign(n1, -1);
    std::vector<bool> vbMatched2(n2, false);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = HISTO_LENGTH / 360.0f;
    int nmatches = 0;
    std::map<unsigned, std::vector<unsigned> >::const_iterator f1it = vFeatVec1.begin(), f2it = vFeatVec2.begin();
    const std::map<unsigned, std::vector<unsigned> >::const_iterator f1end = vFeatVec1.end(), f2end = vFeatVec2.end();
    while (f1it != f1end && f2it != f2end) {
        if (f1it->first == f2it->first) {
            for (size_t i1 = 0, iend1 = f1it->second.size(); i1 < iend1; i1++) {
                const size_t idx1 = f1it->second[i1];
                if (!has_mp1[idx1]) continue;
                const uint8_t* d1 = Descriptors1 + idx1 * 32;
                int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
                for (size_t i2 = 0, iend2 = f2it->second.size(); i2 < iend2; i2++) {
                    const size_t idx2 = f2it->second[i2];
                    if (vbMatched2[idx2] || !has_mp2[idx2]) continue;
                    const int dist = DescriptorDistance(d1, Descriptors2 + idx2 * 32);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = (int)idx2; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (bestDist1 < TH_LOW) {
                    if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
                        vpMatches12[idx1] = bestIdx2;
                        vbMatched2[bestIdx2] = true;
                        if (mbCheckOrientation) {
                            float rot = angle1[idx1] - angle2[bestIdx2];
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)round(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back((int)idx1);
                        }
                        nmatches++;
                    }
                }
            }
            f1it++; f2it++;
        } else if (f1it->first < f2it->first) f1it = vFeatVec1.lower_bound(f2it->first);
        else f2it = vFeatVec2.lower_bound(f1it->first);
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) { vpMatches12[rotHist[i][j]] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// ORBmatcher::CheckDistEpipolarLine (ORBmatcher.cpp:166-190); F12 row-major.
static inline bool CheckDistEpipolarLine(const KeyPoint& kp1, const KeyPoint& kp2, const float* F12, const float* mvLevelSigma2_2)
{
    const float a = kp1.x * F12[0 * 3 + 0] + kp1.y * F12[1 * 3 + 0] + F12[2 * 3 + 0];
    const float b = kp1.x * F12[0 * 3 + 1] + kp1.y * F12[1 * 3 + 1] + F12[2 * 3 + 1];
    const float c = kp1.x * F12[0 * 3 + 2] + kp1.y * F12[1 * 3 + 2] + F12[2 * 3 + 2];
    const float num = a * kp2.x + b * kp2.y + c;
    const float den = a * a + b * b;
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * mvLevelSigma2_2[kp2.octave];
}

// ORBmatcher::SearchForTriangulation (ORBmatcher.cpp:755-927) on flattened arrays; hasMP* = GetMapPoint(i) != NULL.
static inline int SearchForTriangulation(const uint8_t* desc1, const KeyPoint* kps1, const float* mvuRight1, const uint8_t* hasMP1, int N1,
                                         const std::map<unsigned, std::vector<unsigned> >& vFeatVec1,
                                         const uint8_t* desc2, const KeyPoint* kps2, const float* mvuRight2, const uint8_t* hasMP2, int N2,
                                         const std::map<unsigned, std::vector<unsigned> >& vFeatVec2,
                                         const float* F12, float ex, float ey, const float* mvScaleFactors2, const float* mvLevelSigma2_2,
                                         bool bOnlyStereo, bool mbCheckOrientation, std::vector<int>& vMatches12)
{
    int nmatches = 0;
    std::vector<bool> vbMatched2(N2, false);
    vMatches12.assign(N1, -1);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = HISTO_LENGTH / 360.0f;
    std::map<unsigned, std::vector<unsigned> >::const_iterator f1it = vFeatVec1.begin(), f2it = vFeatVec2.begin();
    const std::map<unsigned, std::vector<unsigned> >::const_iterator f1end = vFeatVec1.end(), f2end = vFeatVec2.end();
    while (f1it != f1end && f2it != f2end) {
        if (f1it->first == f2it->first) {
            for (size_t i1 = 0, iend1 = f1it->second.size(); i1 < iend1; i1++) {
                const size_t idx1 = f1it->second[i1];
                if (hasMP1[idx1]) continue;
                const bool bStereo1 = mvuRight1 && mvuRight1[idx1] >= 0;
                if (bOnlyStereo) if (!bStereo1) continue;
                const KeyPoint& kp1 = kps1[idx1];
                const uint8_t* d1 = desc1 + idx1 * 32;
                int bestDist = TH_LOW, bestIdx2 = -1;
                for (size_t i2 = 0, iend2 = f2it->second.size(); i2 < iend2; i2++) {
                    size_t idx2 = f2it->second[i2];
                    if (vbMatched2[idx2] || hasMP2[idx2]) continue;
                    const bool bStereo2 = mvuRight2 && mvuRight2[idx2] >= 0;
                    if (bOnlyStereo) if (!bStereo2) continue;
                    const int dist = DescriptorDistance(d1, desc2 + idx2 * 32);
                    if (dist > TH_LOW || dist > bestDist) continue;
                    const KeyPoint& kp2 = kps2[idx2];
                    if (!bStereo1 && !bStereo2) {
                        const float distex = ex - kp2.x, distey = ey - kp2.y;
                        if (distex * distex + distey * distey < 100 * mvScaleFactors2[kp2.octave]) continue;
                    }
                    if (CheckDistEpipolarLine(kp1, kp2, F12, mvLevelSigma2_2)) { bestIdx2 = (int)idx2; bestDist = dist; }
                }
                if (bestIdx2 >= 0) {
                    const KeyPoint& kp2 = kps2[bestIdx2];
                    vMatches12[idx1] = bestIdx2;
                    vbMatched2[bestIdx2] = true;
                    nmatches++;
                    if (mbCheckOrientation) {
                        float rot = kp1.angle - kp2.angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back((int)idx1);
                    }
                }
            }
            f1it++; f2it++;
        } else if (f1it->first < f2it->first) f1it = vFeatVec1.lower_bound(f2it->first);
        else f2it = vFeatVec2.lower_bound(f1it->first);
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) { vMatches12[rotHist[i][j]] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// Brute-force top-2 with the SearchByBoW update rule (ORBmatcher.cpp:249-275): strict '<' updates,
// first index wins ties, second may equal best.  
static inline void bf_top2(const uint8_t* db, long M, const uint8_t* q, int nq, int32_t* d1, int64_t* i1, int32_t* d2)
{
    for (int i = 0; i < nq; i++) {
        int b1 = 256, b2 = 256; long bi = -1;
        for (long j = 0; j < M; j++) {
            int d = DescriptorDistance(q + (size_t)i * 32, db + (size_t)j * 32);
            if (d < b1) { b2 = b1; b1 = d; bi = j; }
            else if (d < b2) b2 = d;
        }
        d1[i] = b1; i1[i] = bi; d2[i] = b2;
    }
}

// One projected query of the SearchByProjection family.
struct ProjQuery {
    float u, v, r;            // window centre and (already scaled) radius
    int minLevel, maxLevel;   // as passed to GetFeaturesInArea
    float uR;                 // projected right coordinate (a10: mTrackProjXR, a11: u - mbf*invzc)
    float maxErrR;            // stereo gate: skip candidate if |uR - mvuRight[idx]| > maxErrR
    uint8_t desc[32];
    int id;                   // caller's identifier (map point index / last-frame index)
    int hasObs;               // pMP->Observations()>0 once assigned
    float angle;              // LastFrame.mvKeysUn[i].angle for the rotation histogram
};

// SearchByProjection(Frame&, vector<MapPoint*>&, th) (ORBmatcher.cpp:63-155).
// frameMP[idx] = id of the assigned map point or -1; frameMPObs[idx] = its Observations()>0.
static inline int SearchByProjection_LocalPoints(const FrameGrid& F, const float* mvuRight, const std::vector<ProjQuery>& qs,
                                                 float nnratio, std::vector<int>& frameMP, std::vector<int>& frameMPObs)
{
    int nmatches = 0;
    std::vector<int> vIndices;
    for (size_t iMP = 0; iMP < qs.size(); iMP++) {
        const ProjQuery& q = qs[iMP];
        F.GetFeaturesInArea(q.u, q.v, q.r, q.minLevel, q.maxLevel, vIndices);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
            if (frameMP[idx] >= 0 && frameMPObs[idx] > 0) continue;
            if (mvuRight && mvuRight[idx] > 0) {
                const float er = fabs(q.uR - mvuRight[idx]);
                if (er > q.maxErrR) continue;
            }
            const int dist = DescriptorDistance(q.desc, F.desc + (size_t)idx * 32);
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel;
                bestLevel = F.kps[idx].octave; bestIdx = idx;
            } else if (dist < bestDist2) {
                bestLevel2 = F.kps[idx].octave; bestDist2 = dist;
            }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
            frameMP[bestIdx] = q.id; frameMPObs[bestIdx] = q.hasObs;
            nmatches++;
        }
    }
    return nmatches;
}

// SearchByProjection(Frame& Current, const Frame& Last, th, bMono) (ORBmatcher.cpp:1506-1659), after the
// projection step (queries carry u, v, radius and the octave window chosen at :1570-1575).
static inline int SearchByProjection_LastFrame(const FrameGrid& F, const float* mvuRight, const std::vector<ProjQuery>& qs,
                                               bool checkOri, std::vector<int>& frameMP, std::vector<int>& frameMPObs)
{
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = HISTO_LENGTH / 360.0f;
    std::vector<int> vIndices2;
    for (size_t i = 0; i < qs.size(); i++) {
        const ProjQuery& q = qs[i];
        F.GetFeaturesInArea(q.u, q.v, q.r, q.minLevel, q.maxLevel, vIndices2);
        if (vIndices2.empty()) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            if (frameMP[i2] >= 0 && frameMPObs[i2] > 0) continue;
            if (mvuRight && mvuRight[i2] > 0) {
                const float er = fabs(q.uR - mvuRight[i2]);
                if (er > q.maxErrR) continue;
            }
            const int dist = DescriptorDistance(q.desc, F.desc + (size_t)i2 * 32);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {
            frameMP[bestIdx2] = q.id; frameMPObs[bestIdx2] = q.hasObs;
            nmatches++;
            if (checkOri) {
                float rot = q.angle - F.kps[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (size_t j = 0; j < rotHist[i].size(); j++) { frameMP[rotHist[i][j]] = -1; nmatches--; }
    }
    return nmatches;
}

// SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound, th, ORBdist)
// (ORBmatcher.cpp:1661-1789) after the projection step.  frameMP[i] = id of CurrentFrame.mvpMapPoints[i] or -1.
static inline int SearchByProjection_Reloc(const FrameGrid& F, const std::vector<ProjQuery>& qs, int ORBdist, bool checkOri,
                                           std::vector<int>& frameMP)
{
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = HISTO_LENGTH / 360.0f;
    std::vector<int> vIndices2;
    for (size_t i = 0; i < qs.size(); i++) {
        const ProjQuery& q = qs[i];
        F.GetFeaturesInArea(q.u, q.v, q.r, q.minLevel, q.maxLevel, vIndices2);
        if (vIndices2.empty()) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            if (frameMP[i2] >= 0) continue;
            const int dist = DescriptorDistance(q.desc, F.desc + (size_t)i2 * 32);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= ORBdist) {
            frameMP[bestIdx2] = q.id;
            nmatches++;
            if (checkOri) {
                float rot = q.angle - F.kps[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) { frameMP[rotHist[i][j]] = -1; nmatches--; }
    }
    return nmatches;
}

// SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, vector<MapPoint*>& vpMatched, th)
// (ORBmatcher.cpp:352-472) after the projection step: KeyFrame::GetFeaturesInArea(u, v, radius) has no level filter
// (KeyFrame.cpp:694-735), the level test follows in the loop.  q.maxLevel carries nPredictedLevel.
static inline int SearchByProjection_Loop(const FrameGrid& KF, const std::vector<ProjQuery>& qs, std::vector<int>& vpMatched)
{
    int nmatches = 0;
    std::vector<int> vIndices;
    for (size_t iMP = 0; iMP < qs.size(); iMP++) {
        const ProjQuery& q = qs[iMP];
        const int nPredictedLevel = q.maxLevel;
        KF.GetFeaturesInArea(q.u, q.v, q.r, -1, -1, vIndices);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestIdx = -1;
        for (int idx : vIndices) {
            if (vpMatched[idx] >= 0) continue;
            const int kpLevel = KF.kps[idx].octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            const int dist = DescriptorDistance(q.desc, KF.desc + (size_t)idx * 32);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (bestDist <= TH_LOW) { vpMatched[bestIdx] = q.id; nmatches++; }
    }
    return nmatches;
}

// The per-point search of Fuse(KeyFrame*, const vector<MapPoint*>&, th) (ORBmatcher.cpp:1025-1085, gate = true),
// Fuse(KeyFrame*, cv::Mat Scw, ...) (:1195-1225) and SearchBySim3 (:1343-1377, :1423-1457) (gate = false).
static inline void FuseBest(const FrameGrid& KF, const std::vector<ProjQuery>& qs, const float* mvuRight, const float* mvInvLevelSigma2,
                            bool gate, int maxDist, std::vector<int>& bestIdxOut, std::vector<int>& bestDistOut)
{
    bestIdxOut.assign(qs.size(), -1); bestDistOut.assign(qs.size(), -1);
    std::vector<int> vIndices;
    for (size_t i = 0; i < qs.size(); i++) {
        const ProjQuery& q = qs[i];
        const int nPredictedLevel = q.maxLevel;
        const float u = q.u, v = q.v, ur = q.uR;
        KF.GetFeaturesInArea(u, v, q.r, -1, -1, vIndices);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestIdx = -1;
        for (int idx : vIndices) {
            const KeyPoint& kp = KF.kps[idx];
            const int kpLevel = kp.octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            if (gate) {
                if (mvuRight && mvuRight[idx] >= 0) {
                    const float kpx = kp.x, kpy = kp.y, kpr = mvuRight[idx];
                    const float ex = u - kpx, ey = v - kpy, er = ur - kpr;
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * mvInvLevelSigma2[kpLevel] > 7.8) continue;
                } else {
                    const float kpx = kp.x, kpy = kp.y;
                    const float ex = u - kpx, ey = v - kpy;
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * mvInvLevelSigma2[kpLevel] > 5.99) continue;
                }
            }
            const int dist = DescriptorDistance(q.desc, KF.desc + (size_t)idx * 32);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (bestIdx >= 0 && bestDist <= maxDist) { bestIdxOut[i] = bestIdx; bestDistOut[i] = bestDist; }
    }
}

// ORBmatcher::SearchForInitialization (ORBmatcher.cpp:474-590)
static inline int SearchForInitialization(const FrameGrid& F1, const FrameGrid& F2, std::vector<float>& prevMatchedXY,
                                          std::vector<int>& vnMatches12, int windowSize, float nnratio, bool checkOri)
{
    int nmatches = 0;
    vnMatches12.assign(F1.N, -1);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = HISTO_LENGTH / 360.0f;
    std::vector<int> vMatchedDistance(F2.N, INT_MAX), vnMatches21(F2.N, -1);
    std::vector<int> vIndices2;
    for (int i1 = 0; i1 < F1.N; i1++) {
        const KeyPoint& kp1 = F1.kps[i1];
        int level1 = kp1.octave;
        if (level1 > 0) continue;
        F2.GetFeaturesInArea(prevMatchedXY[2 * i1], prevMatchedXY[2 * i1 + 1], (float)windowSize, level1, level1, vIndices2);
        if (vIndices2.empty()) continue;
        const uint8_t* d1 = F1.desc + (size_t)i1 * 32;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            int dist = DescriptorDistance(d1, F2.desc + (size_t)i2 * 32);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * nnratio) {
                if (vnMatches21[bestIdx2] >= 0) { vnMatches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
                vnMatches12[i1] = bestIdx2; vnMatches21[bestIdx2] = i1; vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (checkOri) {
                    float rot = F1.kps[i1].angle - F2.kps[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(i1);
                }
            }
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); j++) {
                int idx1 = rotHist[i][j];
                if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; nmatches--; }
            }
        }
    }
    for (int i1 = 0; i1 < F1.N; i1++)
        if (vnMatches12[i1] >= 0) {
            prevMatchedXY[2 * i1] = F2.kps[vnMatches12[i1]].x;
            prevMatchedXY[2 * i1 + 1] = F2.kps[vnMatches12[i1]].y;
        }
    return nmatches;
}

// Frame::ComputeStereoMatches (Frame.cpp:540-748).  mb is whatever the member holds at the call
// (0 in the stereo constructor => maxD = +inf).
static inline void ComputeStereoMatches(const KeyPoint* keysL, const uint8_t* descL, int N,
                                        const KeyPoint* keysR, const uint8_t* descR, int Nr,
                                        const std::vector<Image>& pyrL, const std::vector<Image>& pyrR,
                                        const std::vector<float>& scaleFactors, const std::vector<float>& invScaleFactors,
                                        float mbf, float mb, float* mvuRight, float* mvDepth,
                                        int* dbgBestIdxR = nullptr, int* dbgBestDist = nullptr)
{
    for (int i = 0; i < N; i++) { mvuRight[i] = -1.0f; mvDepth[i] = -1.0f; if (dbgBestIdxR) { dbgBestIdxR[i] = -1; dbgBestDist[i] = -1; } }
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const int nRows = pyrL[0].h;
    std::vector<std::vector<int>> vRowIndices(nRows);
    for (int iR = 0; iR < Nr; iR++) {
        const KeyPoint& kp = keysR[iR];
        const float kpY = kp.y;
        const float r = 2.0f * scaleFactors[kp.octave];
        const int maxr = (int)ceil(kpY + r), minr = (int)floor(kpY - r);
        for (int yi = minr; yi <= maxr; yi++)
            if (yi >= 0 && yi < nRows) vRowIndices[yi].push_back(iR);   // guard: the reference indexes unchecked
    }
    const float minZ = mb, minD = 0, maxD = mbf / minZ;
    std::vector<std::pair<int, int>> vDistIdx;
    for (int iL = 0; iL < N; iL++) {
        const KeyPoint& kpL = keysL[iL];
        const int levelL = kpL.octave;
        const float vL = kpL.y, uL = kpL.x;
        const std::vector<int>& vCandidates = vRowIndices[(int)vL];
        if (vCandidates.empty()) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH; int bestIdxR = 0;
        const uint8_t* dL = descL + (size_t)iL * 32;
        for (size_t iC = 0; iC < vCandidates.size(); iC++) {
            const int iR = vCandidates[iC];
            const KeyPoint& kpR = keysR[iR];
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = DescriptorDistance(dL, descR + (size_t)iR * 32);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < thOrbDist) {
            if (dbgBestIdxR) { dbgBestIdxR[iL] = bestIdxR; dbgBestDist[iL] = bestDist; }
            const float uR0 = keysR[bestIdxR].x;
            const float scaleFactor = invScaleFactors[kpL.octave];
            const float scaleduL = round(kpL.x * scaleFactor);
            const float scaledvL = round(kpL.y * scaleFactor);
            const float scaleduR0 = round(uR0 * scaleFactor);
            const int w = 5;
            const Image& imL = pyrL[kpL.octave];
            const Image& imR = pyrR[kpL.octave];
            const int cxL = (int)scaleduL, cyL = (int)scaledvL;
            int sadBest = INT_MAX, bestincR = 0;
            const int L = 5;
            std::vector<float> vDists(2 * L + 1);
            const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= imR.w) continue;
            for (int incR = -L; incR <= +L; incR++) {
                const int cxR = (int)scaleduR0 + incR;
                const float cL = (float)imL.row(cyL)[cxL], cR = (float)imR.row(cyL)[cxR];
                double acc = 0;   // cv::norm(IL, IR, NORM_L1) on CV_32F windows (integers: exact)
                for (int dy = -w; dy <= w; dy++)
                    for (int dx = -w; dx <= w; dx++) {
                        float a = (float)imL.row(cyL + dy)[cxL + dx] - cL;
                        float b = (float)imR.row(cyL + dy)[cxR + dx] - cR;
                        acc += fabs(a - b);
                    }
                float dist = (float)acc;
                if (dist < sadBest) { sadBest = (int)dist; bestincR = incR; }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;
            const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = scaleFactors[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) { disparity = (float)0.01; bestuR = (float)(uL - 0.01); }   // double literals as in :716-717
                mvDepth[iL] = mbf / disparity;
                mvuRight[iL] = bestuR;
                vDistIdx.push_back(std::pair<int, int>(sadBest, iL));
            }
        }
    }
    if (vDistIdx.empty()) return;   // the reference indexes vDistIdx[0] unchecked here
    std::sort(vDistIdx.begin(), vDistIdx.end());
    const float median = (float)vDistIdx[vDistIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
        if (vDistIdx[i].first < thDist) break;
        mvuRight[vDistIdx[i].second] = -1;
        mvDepth[vDistIdx[i].second] = -1;
    }
}

}  // namespace orbo
