// oracle/refmatch_capi.cpp — TEST INFRASTRUCTURE.  C API (ctypes) over the reference's OWN matching code: src/ORBmatcher.cpp,
// src/Frame.cpp, src/KeyFrame.cpp, src/MapPoint.cpp, src/Map.cpp, src/KeyFrameDatabase.cpp, src/ORBextractor.cpp and the vendored
// Thirdparty/DBoW2 (BowVector, FeatureVector, FORB, ScoringObject, TemplatedVocabulary.h, DUtils/Random), all compiled UNMODIFIED
// from /root/reference/ORB-SLAM2 over oracle/cvshim (recipe: oracle/Makefile, target `refmatch`; output oracle/_ref/liborb_refmatch.so).
// Nothing below restates an algorithm: it builds Frame / KeyFrame / MapPoint objects from arrays (compiled with
// -fno-access-control so protected members can be filled), calls the reference's member functions and copies their results out.
//
// Two facilities make the reference's INTERNAL values observable without touching its text:
//   * the linker's --wrap on Frame::GetFeaturesInArea, KeyFrame::GetFeaturesInArea and KeyFrame::GetMapPoint records every call
//     ORBmatcher.cpp makes (window centre, radius, level bounds; index asked for), in order — that is how the fixtures learn the
//     exact projected queries and the per-point "bestIdx" of Fuse;
//   * a monotonic allocator during Frame construction makes ORBextractor.cpp:690's address tie-break the canonical
//     creation-order one (same device as oracle/ref_capi.cpp).
#include <new>
#include <cstdio>
#include <atomic>
#include <map>
#include <set>
#include <sys/mman.h>
#include "ORBmatcher.h"
#include "Frame.h"
#include "KeyFrame.h"
#include "MapPoint.h"
#include "Map.h"
#include "ORBVocabulary.h"
#include "Converter.h"

using namespace ORB_SLAM2;

// ---- allocator: malloc normally, monotonic bump while g_bump_on (addresses grow with creation order) -------------------------
namespace {
const size_t ARENA_BYTES = size_t(1) << 36;          // 64 GiB of address space, committed lazily, never rewound
char* g_arena = nullptr;
std::atomic<size_t> g_used(0);
std::atomic<int> g_bump_on(0);
inline void arena_init()
{
    if (!g_arena) {
        g_arena = (char*)mmap(nullptr, ARENA_BYTES, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (g_arena == (char*)MAP_FAILED) { fprintf(stderr, "refmatch: arena mmap failed\n"); abort(); }
    }
}
inline void* take(size_t n)
{
    if (g_bump_on.load(std::memory_order_relaxed)) {
        n = (n + 15) & ~size_t(15);
        const size_t off = g_used.fetch_add(n);
        if (off + n > ARENA_BYTES) { fprintf(stderr, "refmatch: arena exhausted\n"); abort(); }
        return g_arena + off;
    }
    void* p = malloc(n ? n : 1);
    if (!p) abort();
    return p;
}
inline void give(void* p)
{
    if (!p) return;
    if (g_arena && (char*)p >= g_arena && (char*)p < g_arena + ARENA_BYTES) return;
    free(p);
}
struct BumpScope { BumpScope() { arena_init(); g_bump_on = 1; } ~BumpScope() { g_bump_on = 0; } };
}
void* operator new(size_t n) { return take(n); }
void* operator new[](size_t n) { return take(n); }
void operator delete(void* p) noexcept { give(p); }
void operator delete[](void* p) noexcept { give(p); }
void operator delete(void* p, size_t) noexcept { give(p); }
void operator delete[](void* p, size_t) noexcept { give(p); }

// src/Converter.cpp:28-37 (the real header needs Eigen / g2o; see oracle/refshim/Converter.h)
std::vector<cv::Mat> Converter::toDescriptorVector(const cv::Mat& Descriptors)
{
    std::vector<cv::Mat> vDesc;
    vDesc.reserve(Descriptors.rows);
    for (int j = 0; j < Descriptors.rows; j++) vDesc.push_back(Descriptors.row(j));
    return vDesc;
}

// ---- call trace (linker --wrap) --------------------------------------------------------------------------------------------
namespace {
struct TraceEvent { int kind; float x, y, r; int a, b; const void* obj; };   // kind 0 Frame::GFA, 1 KeyFrame::GFA, 2 KeyFrame::GetMapPoint, 3 MapPoint::GetWorldPos, 4 MapPoint::isBad, 5 MapPoint::PredictScale
std::vector<TraceEvent> g_trace;
bool g_trace_on = false;
}
extern "C" {
std::vector<size_t> __real__ZNK9ORB_SLAM25Frame17GetFeaturesInAreaERKfS2_S2_ii(const Frame*, const float&, const float&, const float&, int, int);
std::vector<size_t> __wrap__ZNK9ORB_SLAM25Frame17GetFeaturesInAreaERKfS2_S2_ii(const Frame* self, const float& x, const float& y, const float& r, int minLevel, int maxLevel)
{
    if (g_trace_on) g_trace.push_back(TraceEvent{0, x, y, r, minLevel, maxLevel, self});
    return __real__ZNK9ORB_SLAM25Frame17GetFeaturesInAreaERKfS2_S2_ii(self, x, y, r, minLevel, maxLevel);
}
std::vector<size_t> __real__ZNK9ORB_SLAM28KeyFrame17GetFeaturesInAreaERKfS2_S2_(const KeyFrame*, const float&, const float&, const float&);
std::vector<size_t> __wrap__ZNK9ORB_SLAM28KeyFrame17GetFeaturesInAreaERKfS2_S2_(const KeyFrame* self, const float& x, const float& y, const float& r)
{
    if (g_trace_on) g_trace.push_back(TraceEvent{1, x, y, r, -1, -1, self});
    return __real__ZNK9ORB_SLAM28KeyFrame17GetFeaturesInAreaERKfS2_S2_(self, x, y, r);
}
MapPoint* __real__ZN9ORB_SLAM28KeyFrame11GetMapPointERKm(KeyFrame*, const size_t&);
MapPoint* __wrap__ZN9ORB_SLAM28KeyFrame11GetMapPointERKm(KeyFrame* self, const size_t& idx)
{
    if (g_trace_on) g_trace.push_back(TraceEvent{2, 0, 0, 0, (int)idx, 0, self});
    return __real__ZN9ORB_SLAM28KeyFrame11GetMapPointERKm(self, idx);
}
// MapPoint::GetWorldPos / isBad: called once per candidate point before its projection, so the event that precedes a
// GetFeaturesInArea event names the MapPoint the window belongs to (kind 3 / 4, obj = the MapPoint).
cv::Mat __real__ZN9ORB_SLAM28MapPoint11GetWorldPosEv(MapPoint*);
cv::Mat __wrap__ZN9ORB_SLAM28MapPoint11GetWorldPosEv(MapPoint* self)
{
    if (g_trace_on) g_trace.push_back(TraceEvent{3, 0, 0, 0, 0, 0, self});
    return __real__ZN9ORB_SLAM28MapPoint11GetWorldPosEv(self);
}
// MapPoint::PredictScale: kind 5, x = the distance passed in, a = the level returned (the [p-1, p] gate of the KeyFrame variants
// is applied inside ORBmatcher.cpp's candidate loop and is otherwise invisible)
int __real__ZN9ORB_SLAM28MapPoint12PredictScaleERKfPNS_5FrameE(MapPoint*, const float&, Frame*);
int __wrap__ZN9ORB_SLAM28MapPoint12PredictScaleERKfPNS_5FrameE(MapPoint* self, const float& d, Frame* f)
{
    const int r = __real__ZN9ORB_SLAM28MapPoint12PredictScaleERKfPNS_5FrameE(self, d, f);
    if (g_trace_on) g_trace.push_back(TraceEvent{5, d, 0, 0, r, 0, self});
    return r;
}
int __real__ZN9ORB_SLAM28MapPoint12PredictScaleERKfPNS_8KeyFrameE(MapPoint*, const float&, KeyFrame*);
int __wrap__ZN9ORB_SLAM28MapPoint12PredictScaleERKfPNS_8KeyFrameE(MapPoint* self, const float& d, KeyFrame* f)
{
    const int r = __real__ZN9ORB_SLAM28MapPoint12PredictScaleERKfPNS_8KeyFrameE(self, d, f);
    if (g_trace_on) g_trace.push_back(TraceEvent{5, d, 0, 0, r, 0, self});
    return r;
}
bool __real__ZN9ORB_SLAM28MapPoint5isBadEv(MapPoint*);
bool __wrap__ZN9ORB_SLAM28MapPoint5isBadEv(MapPoint* self)
{
    if (g_trace_on) g_trace.push_back(TraceEvent{4, 0, 0, 0, 0, 0, self});
    return __real__ZN9ORB_SLAM28MapPoint5isBadEv(self);
}
}

namespace {
struct MapPointSet {
    std::vector<MapPoint*> v;
    std::map<MapPoint*, int> index;
    int indexOf(MapPoint* p) const { if (!p) return -1; auto it = index.find(p); return it == index.end() ? -2 : it->second; }
};
cv::Mat mat32f(int r, int c, const float* v)
{
    cv::Mat m(r, c, CV_32F);
    for (int i = 0; i < r; i++) for (int j = 0; j < c; j++) m.at<float>(i, j) = v[i * c + j];
    return m;
}
cv::Mat matK(const float* K4)
{
    cv::Mat K = cv::Mat::eye(3, 3, CV_32F);
    K.at<float>(0, 0) = K4[0]; K.at<float>(1, 1) = K4[1]; K.at<float>(0, 2) = K4[2]; K.at<float>(1, 2) = K4[3];
    return K;
}
}

extern "C" {

// ---- bookkeeping ----------------------------------------------------------------------------------------------------------
void rm_reset()
{
    Frame::mbInitialComputations = true;     // the next Frame recomputes image bounds, grid pitch and fx .. invfy (Frame.cpp:95-111)
    Frame::nNextId = 0; KeyFrame::nNextId = 0; MapPoint::nNextId = 0;
    g_trace.clear();
}
void rm_trace(int on) { g_trace_on = on != 0; if (on) g_trace.clear(); }
int rm_trace_count() { return (int)g_trace.size(); }
// kind[n], xyr[3n], ab[2n], obj[n]
void rm_trace_get(int32_t* kind, float* xyr, int32_t* ab, uint64_t* obj)
{
    for (size_t i = 0; i < g_trace.size(); i++) {
        kind[i] = g_trace[i].kind; xyr[3 * i] = g_trace[i].x; xyr[3 * i + 1] = g_trace[i].y; xyr[3 * i + 2] = g_trace[i].r;
        ab[2 * i] = g_trace[i].a; ab[2 * i + 1] = g_trace[i].b; obj[i] = (uint64_t)(uintptr_t)g_trace[i].obj;
    }
}

// ---- vocabulary (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h) ------------------------------------------------------------------
void* rm_vocab_load(const char* path)
{
    ORBVocabulary* v = new ORBVocabulary();
    if (!v->loadFromTextFile(path)) { delete v; return nullptr; }
    return v;
}
int rm_vocab_size(void* voc) { return (int)((ORBVocabulary*)voc)->size(); }
// TemplatedVocabulary::transform(feature, id, weight, nid, levelsup) (:1242-1285), one descriptor at a time
void rm_vocab_transform(void* voc, const uint8_t* desc, int n, int levelsup, int32_t* word_id, double* weight, int32_t* node_id)
{
    ORBVocabulary* v = (ORBVocabulary*)voc;
    for (int i = 0; i < n; i++) {
        cv::Mat d(1, 32, CV_8U, (void*)(desc + (size_t)i * 32));
        DBoW2::WordId id = 0; DBoW2::WordValue w = 0; DBoW2::NodeId nid = 0;
        v->transform(d, id, w, &nid, levelsup);
        word_id[i] = (int32_t)id; weight[i] = (double)w; node_id[i] = (int32_t)nid;
    }
}
static void dump_bow(const DBoW2::BowVector& bv, const DBoW2::FeatureVector& fv, int32_t* bow_ids, double* bow_vals, int32_t* n_bow,
                     int32_t* fv_nodes, int32_t* fv_off, int32_t* fv_feat, int32_t* n_fv)
{
    int k = 0;
    for (auto it = bv.begin(); it != bv.end(); ++it, ++k) { bow_ids[k] = (int32_t)it->first; bow_vals[k] = it->second; }
    *n_bow = k;
    int j = 0, o = 0;
    for (auto it = fv.begin(); it != fv.end(); ++it, ++j) {
        fv_nodes[j] = (int32_t)it->first; fv_off[j] = o;
        for (size_t t = 0; t < it->second.size(); t++) fv_feat[o++] = (int32_t)it->second[t];
    }
    fv_off[j] = o; *n_fv = j;
}
// transform(features, BowVector&, FeatureVector&, levelsup) (:1151-1218)
void rm_vocab_bow(void* voc, const uint8_t* desc, int n, int levelsup, int32_t* bow_ids, double* bow_vals, int32_t* n_bow,
                  int32_t* fv_nodes, int32_t* fv_off, int32_t* fv_feat, int32_t* n_fv)
{
    cv::Mat D(n, 32, CV_8U, (void*)desc);
    std::vector<cv::Mat> vd = Converter::toDescriptorVector(D);
    DBoW2::BowVector bv; DBoW2::FeatureVector fv;
    ((ORBVocabulary*)voc)->transform(vd, bv, fv, levelsup);
    dump_bow(bv, fv, bow_ids, bow_vals, n_bow, fv_nodes, fv_off, fv_feat, n_fv);
}

// ---- extractor + frames ---------------------------------------------------------------------------------------------------
void* rm_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh)
{
    return new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
}
int rm_extractor_level(void* ex, int level, uint8_t* dst, int* w, int* h)
{
    const cv::Mat& m = ((ORBextractor*)ex)->mvImagePyramid[level];
    *w = m.cols; *h = m.rows;
    if (dst) for (int y = 0; y < m.rows; y++) memcpy(dst + (size_t)y * m.cols, m.ptr(y), m.cols);
    return 0;
}
// Frame(imGray, timeStamp, extractor, voc, K, distCoef, bf, thDepth) (Frame.cpp:186-235): extraction, UndistortKeyPoints,
// ComputeImageBounds (first frame after rm_reset), AssignFeaturesToGrid — all the reference's.
void* rm_frame_mono(const uint8_t* img, int w, int h, void* ex, void* voc, const float* K4, const float* dist, int nd, float bf, float thDepth)
{
    BumpScope bump;
    cv::Mat im(h, w, CV_8UC1, (void*)img, (size_t)w), K = matK(K4), D = mat32f(nd, 1, dist);
    return new Frame(im, 0.0, (ORBextractor*)ex, (ORBVocabulary*)voc, K, D, bf, thDepth);
}
// Frame(imLeft, imRight, ...) (Frame.cpp:79-130): two extraction threads, UndistortKeyPoints, ComputeStereoMatches, grid.
void* rm_frame_stereo(const uint8_t* imgL, const uint8_t* imgR, int w, int h, void* exL, void* exR, void* voc, const float* K4,
                      const float* dist, int nd, float bf, float thDepth)
{
    BumpScope bump;
    cv::Mat l(h, w, CV_8UC1, (void*)imgL, (size_t)w), r(h, w, CV_8UC1, (void*)imgR, (size_t)w), K = matK(K4), D = mat32f(nd, 1, dist);
    return new Frame(l, r, 0.0, (ORBextractor*)exL, (ORBextractor*)exR, (ORBVocabulary*)voc, K, D, bf, thDepth);
}
void* rm_frame_copy(void* f) { return new Frame(*(Frame*)f); }
int rm_frame_n(void* f) { return ((Frame*)f)->N; }
int rm_frame_n_right(void* f) { return (int)((Frame*)f)->mvKeysRight.size(); }
void rm_frame_get(void* fp, void* kps, void* kpsUn, uint8_t* desc, float* uRight, float* depth, void* kpsRight, uint8_t* descRight)
{
    Frame* f = (Frame*)fp;
    static_assert(sizeof(cv::KeyPoint) == 28, "cv::KeyPoint layout");
    if (kps) memcpy(kps, f->mvKeys.data(), f->mvKeys.size() * 28);
    if (kpsUn) memcpy(kpsUn, f->mvKeysUn.data(), f->mvKeysUn.size() * 28);
    if (desc) for (int i = 0; i < f->mDescriptors.rows; i++) memcpy(desc + (size_t)i * 32, f->mDescriptors.ptr(i), 32);
    if (uRight && !f->mvuRight.empty()) memcpy(uRight, f->mvuRight.data(), f->mvuRight.size() * 4);
    if (depth && !f->mvDepth.empty()) memcpy(depth, f->mvDepth.data(), f->mvDepth.size() * 4);
    if (kpsRight) memcpy(kpsRight, f->mvKeysRight.data(), f->mvKeysRight.size() * 28);
    if (descRight) for (int i = 0; i < f->mDescriptorsRight.rows; i++) memcpy(descRight + (size_t)i * 32, f->mDescriptorsRight.ptr(i), 32);
}
// {fx, fy, cx, cy, invfx, invfy, mnMinX, mnMinY, mnMaxX, mnMaxY, mfGridElementWidthInv, mfGridElementHeightInv}
void rm_frame_statics(float* o)
{
    o[0] = Frame::fx; o[1] = Frame::fy; o[2] = Frame::cx; o[3] = Frame::cy; o[4] = Frame::invfx; o[5] = Frame::invfy;
    o[6] = Frame::mnMinX; o[7] = Frame::mnMinY; o[8] = Frame::mnMaxX; o[9] = Frame::mnMaxY;
    o[10] = Frame::mfGridElementWidthInv; o[11] = Frame::mfGridElementHeightInv;
}
// scale tables (nlevels entries each), {mbf, mb, mThDepth, mfScaleFactor, mfLogScaleFactor}
void rm_frame_scales(void* fp, float* scale, float* invScale, float* sigma2, float* invSigma2, float* misc5, int32_t* nlevels)
{
    Frame* f = (Frame*)fp;
    *nlevels = f->mnScaleLevels;
    for (int i = 0; i < f->mnScaleLevels; i++) {
        scale[i] = f->mvScaleFactors[i]; invScale[i] = f->mvInvScaleFactors[i]; sigma2[i] = f->mvLevelSigma2[i]; invSigma2[i] = f->mvInvLevelSigma2[i];
    }
    misc5[0] = f->mbf; misc5[1] = f->mb; misc5[2] = f->mThDepth; misc5[3] = f->mfScaleFactor; misc5[4] = f->mfLogScaleFactor;
}
void rm_frame_set_pose(void* fp, const float* Tcw16) { ((Frame*)fp)->SetPose(mat32f(4, 4, Tcw16)); }
// Frame::GetFeaturesInArea (Frame.cpp:379-432), called directly (not traced)
int rm_frame_features_in_area(void* fp, float x, float y, float r, int minLevel, int maxLevel, int32_t* out, int cap)
{
    std::vector<size_t> v = __real__ZNK9ORB_SLAM25Frame17GetFeaturesInAreaERKfS2_S2_ii((Frame*)fp, x, y, r, minLevel, maxLevel);
    for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = (int32_t)v[i];
    return (int)v.size();
}
// mGrid[64][48] dumped as CSR in (col, row) order: off[64*48+1], idx[N]
void rm_frame_grid(void* fp, int32_t* off, int32_t* idx)
{
    Frame* f = (Frame*)fp;
    int o = 0;
    for (int i = 0; i < FRAME_GRID_COLS; i++)
        for (int j = 0; j < FRAME_GRID_ROWS; j++) {
            off[i * FRAME_GRID_ROWS + j] = o;
            for (size_t t = 0; t < f->mGrid[i][j].size(); t++) idx[o++] = (int32_t)f->mGrid[i][j][t];
        }
    off[FRAME_GRID_COLS * FRAME_GRID_ROWS] = o;
}
// Frame::ComputeBoW (Frame.cpp:452-459)
void rm_frame_compute_bow(void* fp, int32_t* bow_ids, double* bow_vals, int32_t* n_bow, int32_t* fv_nodes, int32_t* fv_off, int32_t* fv_feat, int32_t* n_fv)
{
    Frame* f = (Frame*)fp;
    f->ComputeBoW();
    if (bow_ids) dump_bow(f->mBowVec, f->mFeatVec, bow_ids, bow_vals, n_bow, fv_nodes, fv_off, fv_feat, n_fv);
}

// ---- map, key frames, map points ---------------------------------------------------------------------------------------------
void* rm_map_create() { return new Map(); }
// KeyFrame(Frame&, Map*, KeyFrameDatabase*) (KeyFrame.cpp:33-61): copies keypoints, descriptors, grid, BoW, MapPoint slots, pose.
void* rm_kf_create(void* frame, void* map) { return new KeyFrame(*(Frame*)frame, (Map*)map, nullptr); }
void rm_kf_set_pose(void* kf, const float* Tcw16) { ((KeyFrame*)kf)->SetPose(mat32f(4, 4, Tcw16)); }
int rm_kf_n(void* kf) { return ((KeyFrame*)kf)->N; }
int rm_kf_features_in_area(void* kf, float x, float y, float r, int32_t* out, int cap)
{
    std::vector<size_t> v = __real__ZNK9ORB_SLAM28KeyFrame17GetFeaturesInAreaERKfS2_S2_((KeyFrame*)kf, x, y, r);
    for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = (int32_t)v[i];
    return (int)v.size();
}
// KeyFrame::ComputeBoW (KeyFrame.cpp:63-75)
void rm_kf_compute_bow(void* kfp, int32_t* bow_ids, double* bow_vals, int32_t* n_bow, int32_t* fv_nodes, int32_t* fv_off, int32_t* fv_feat, int32_t* n_fv)
{
    KeyFrame* kf = (KeyFrame*)kfp;
    kf->ComputeBoW();
    if (bow_ids) dump_bow(kf->mBowVec, kf->mFeatVec, bow_ids, bow_vals, n_bow, fv_nodes, fv_off, fv_feat, n_fv);
}
// {GetRotation (9), GetTranslation (3), GetCameraCenter (3)}
void rm_kf_pose(void* kfp, float* R9, float* t3, float* O3)
{
    KeyFrame* kf = (KeyFrame*)kfp;
    cv::Mat R = kf->GetRotation(), t = kf->GetTranslation(), O = kf->GetCameraCenter();
    for (int i = 0; i < 3; i++) { for (int j = 0; j < 3; j++) R9[3 * i + j] = R.at<float>(i, j); t3[i] = t.at<float>(i); O3[i] = O.at<float>(i); }
}

// n MapPoints (MapPoint(Pos, pRefKF, pMap), MapPoint.cpp:41-56) with their protected state filled from arrays: representative
// descriptor (what ComputeDistinctiveDescriptors would have stored), mean viewing direction, scale-invariance distances, and nObs.
void* rm_mappoints_create(void* map, void* refkf, int n, const float* pos, const uint8_t* desc, const float* normal,
                          const float* minDist, const float* maxDist, const int32_t* nobs)
{
    MapPointSet* s = new MapPointSet();
    for (int i = 0; i < n; i++) {
        MapPoint* p = new MapPoint(mat32f(3, 1, pos + 3 * i), (KeyFrame*)refkf, (Map*)map);
        cv::Mat d(1, 32, CV_8U, (void*)(desc + (size_t)i * 32));
        p->mDescriptor = d.clone();
        if (normal) p->mNormalVector = mat32f(3, 1, normal + 3 * i);
        if (minDist) p->mfMinDistance = minDist[i];
        if (maxDist) p->mfMaxDistance = maxDist[i];
        if (nobs) p->nObs = nobs[i];
        s->index[p] = i;
        s->v.push_back(p);
    }
    return s;
}
void rm_mappoints_ptrs(void* s, uint64_t* out) { MapPointSet* m = (MapPointSet*)s; for (size_t i = 0; i < m->v.size(); i++) out[i] = (uint64_t)(uintptr_t)m->v[i]; }
int rm_mappoints_n(void* s) { return (int)((MapPointSet*)s)->v.size(); }
// what Frame::isInFrustum (Frame.cpp:305-372) leaves on each MapPoint for SearchByProjection(F, vpMapPoints, th)
void rm_mappoints_set_track(void* sp, const uint8_t* inView, const float* projX, const float* projY, const float* projXR,
                            const int32_t* level, const float* viewCos)
{
    MapPointSet* s = (MapPointSet*)sp;
    for (size_t i = 0; i < s->v.size(); i++) {
        MapPoint* p = s->v[i];
        p->mbTrackInView = inView[i] != 0; p->mTrackProjX = projX[i]; p->mTrackProjY = projY[i]; p->mTrackProjXR = projXR[i];
        p->mnTrackScaleLevel = level[i]; p->mTrackViewCos = viewCos[i];
    }
}
// Frame::isInFrustum itself (Frame.cpp:305-372): fills the tracking fields from the frame's pose; returns the flags
void rm_mappoints_in_frustum(void* sp, void* frame, float viewingCosLimit, uint8_t* inView, float* projX, float* projY, float* projXR,
                             int32_t* level, float* viewCos)
{
    MapPointSet* s = (MapPointSet*)sp;
    for (size_t i = 0; i < s->v.size(); i++) {
        MapPoint* p = s->v[i];
        inView[i] = ((Frame*)frame)->isInFrustum(p, viewingCosLimit) ? 1 : 0;
        projX[i] = p->mTrackProjX; projY[i] = p->mTrackProjY; projXR[i] = p->mTrackProjXR; level[i] = p->mnTrackScaleLevel; viewCos[i] = p->mTrackViewCos;
    }
}
void rm_mappoints_set_bad(void* sp, int i) { ((MapPointSet*)sp)->v[i]->mbBad = true; }
void rm_mappoints_add_observation(void* sp, int i, void* kf, int idx) { ((MapPointSet*)sp)->v[i]->AddObservation((KeyFrame*)kf, (size_t)idx); }
void rm_mappoints_nobs(void* sp, int32_t* out) { MapPointSet* s = (MapPointSet*)sp; for (size_t i = 0; i < s->v.size(); i++) out[i] = s->v[i]->Observations(); }
void rm_mappoints_is_bad(void* sp, uint8_t* out) { MapPointSet* s = (MapPointSet*)sp; for (size_t i = 0; i < s->v.size(); i++) out[i] = s->v[i]->isBad() ? 1 : 0; }
// MapPoint::PredictScale(dist, Frame*) / (dist, KeyFrame*) (MapPoint.cpp:481-520)
void rm_mappoints_predict_scale(void* sp, const float* dist, void* frame, void* kf, int32_t* out)
{
    MapPointSet* s = (MapPointSet*)sp;
    for (size_t i = 0; i < s->v.size(); i++) out[i] = frame ? s->v[i]->PredictScale(dist[i], (Frame*)frame) : s->v[i]->PredictScale(dist[i], (KeyFrame*)kf);
}
// MapPoint::ComputeDistinctiveDescriptors (MapPoint.cpp:302-380) on point i with its real observations; returns GetDescriptor()
void rm_mappoints_distinctive(void* sp, int i, uint8_t* desc32)
{
    MapPoint* p = ((MapPointSet*)sp)->v[i];
    p->ComputeDistinctiveDescriptors();
    cv::Mat d = p->GetDescriptor();
    memcpy(desc32, d.ptr(0), 32);
}
// observation order of point i (std::map<KeyFrame*, size_t> iteration order = what ComputeDistinctiveDescriptors walks): kf handles + idx
int rm_mappoints_observations(void* sp, int i, uint64_t* kfs, int32_t* idx, int cap)
{
    std::map<KeyFrame*, size_t> obs = ((MapPointSet*)sp)->v[i]->GetObservations();
    int k = 0;
    for (auto it = obs.begin(); it != obs.end() && k < cap; ++it, ++k) { kfs[k] = (uint64_t)(uintptr_t)it->first; idx[k] = (int32_t)it->second; }
    return (int)obs.size();
}

// Frame.mvpMapPoints / mvbOutlier from an index array into a MapPointSet (-1 = NULL)
void rm_frame_set_mappoints(void* fp, void* sp, const int32_t* idx, const uint8_t* outlier)
{
    Frame* f = (Frame*)fp; MapPointSet* s = (MapPointSet*)sp;
    f->mvpMapPoints.assign(f->N, (MapPoint*)nullptr);
    f->mvbOutlier.assign(f->N, false);
    for (int i = 0; i < f->N; i++) {
        if (idx && idx[i] >= 0) f->mvpMapPoints[i] = s->v[idx[i]];
        if (outlier) f->mvbOutlier[i] = outlier[i] != 0;
    }
}
void rm_frame_get_mappoints(void* fp, void* sp, int32_t* idx)
{
    Frame* f = (Frame*)fp; MapPointSet* s = (MapPointSet*)sp;
    for (int i = 0; i < f->N; i++) idx[i] = s->indexOf(f->mvpMapPoints[i]);
}
// KeyFrame slots: AddMapPoint + (optionally) the MapPoint's AddObservation, as Tracking / LocalMapping pair them
void rm_kf_set_mappoints(void* kfp, void* sp, const int32_t* idx, int add_observation)
{
    KeyFrame* kf = (KeyFrame*)kfp; MapPointSet* s = (MapPointSet*)sp;
    for (int i = 0; i < kf->N; i++) {
        if (idx[i] < 0) continue;
        kf->AddMapPoint(s->v[idx[i]], (size_t)i);
        if (add_observation) s->v[idx[i]]->AddObservation(kf, (size_t)i);
    }
}
void rm_kf_get_mappoints(void* kfp, void* sp, int32_t* idx)
{
    KeyFrame* kf = (KeyFrame*)kfp; MapPointSet* s = (MapPointSet*)sp;
    std::vector<MapPoint*> v = kf->GetMapPointMatches();
    for (int i = 0; i < kf->N; i++) idx[i] = s->indexOf(v[i]);
}

// ---- the reference's search functions ----------------------------------------------------------------------------------------
// ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th) (ORBmatcher.cpp:63-155)
int rm_search_local_points(void* frame, void* sp, float th, float nnratio)
{
    ORBmatcher m(nnratio, true);
    return m.SearchByProjection(*(Frame*)frame, ((MapPointSet*)sp)->v, th);
}
// ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th, bMono) (:1506-1659)
int rm_search_last_frame(void* cur, void* last, float th, int mono, float nnratio, int checkOri)
{
    ORBmatcher m(nnratio, checkOri != 0);
    return m.SearchByProjection(*(Frame*)cur, *(const Frame*)last, th, mono != 0);
}
// ORBmatcher::SearchByProjection(Frame& Current, KeyFrame*, const set<MapPoint*>& sAlreadyFound, th, ORBdist) (:1661-1789)
int rm_search_reloc(void* cur, void* kf, void* sp, const int32_t* already, int nalready, float th, int ORBdist, float nnratio, int checkOri)
{
    MapPointSet* s = (MapPointSet*)sp;
    std::set<MapPoint*> found;
    for (int i = 0; i < nalready; i++) found.insert(s->v[already[i]]);
    ORBmatcher m(nnratio, checkOri != 0);
    return m.SearchByProjection(*(Frame*)cur, (KeyFrame*)kf, found, th, ORBdist);
}
// ORBmatcher::SearchByProjection(KeyFrame*, cv::Mat Scw, const vector<MapPoint*>& vpPoints, vector<MapPoint*>& vpMatched, th) (:352-472)
// matched (in/out): MapPointSet index per KeyFrame feature, -1 = NULL
int rm_search_loop(void* kfp, const float* Scw16, void* sp, int32_t* matched, int th, float nnratio)
{
    KeyFrame* kf = (KeyFrame*)kfp; MapPointSet* s = (MapPointSet*)sp;
    std::vector<MapPoint*> vm(kf->N, (MapPoint*)nullptr);
    for (int i = 0; i < kf->N; i++) if (matched[i] >= 0) vm[i] = s->v[matched[i]];
    ORBmatcher m(nnratio, true);
    const int n = m.SearchByProjection(kf, mat32f(4, 4, Scw16), s->v, vm, th);
    for (int i = 0; i < kf->N; i++) matched[i] = s->indexOf(vm[i]);
    return n;
}
// ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&) (:205-348): out[iF] = MapPointSet index or -1
int rm_search_by_bow_kf_frame(void* kf, void* fp, void* sp, int32_t* out, float nnratio, int checkOri)
{
    Frame* f = (Frame*)fp;
    std::vector<MapPoint*> vm;
    ORBmatcher m(nnratio, checkOri != 0);
    const int n = m.SearchByBoW((KeyFrame*)kf, *f, vm);
    for (int i = 0; i < f->N; i++) out[i] = ((MapPointSet*)sp)->indexOf(vm[i]);
    return n;
}
// ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&) (:604-743): out[i1] = MapPointSet index (a point of KF2) or -1
int rm_search_by_bow_kf_kf(void* kf1, void* kf2, void* sp, int32_t* out, float nnratio, int checkOri)
{
    std::vector<MapPoint*> vm;
    ORBmatcher m(nnratio, checkOri != 0);
    const int n = m.SearchByBoW((KeyFrame*)kf1, (KeyFrame*)kf2, vm);
    for (size_t i = 0; i < vm.size(); i++) out[i] = ((MapPointSet*)sp)->indexOf(vm[i]);
    return n;
}
// ORBmatcher::SearchForInitialization (:474-590): prevMatched (in/out, 2*N1 floats), matches12 out
int rm_search_for_initialization(void* f1p, void* f2p, float* prevMatched, int32_t* matches12, int windowSize, float nnratio, int checkOri)
{
    Frame* f1 = (Frame*)f1p;
    std::vector<cv::Point2f> prev(f1->N);
    for (int i = 0; i < f1->N; i++) prev[i] = cv::Point2f(prevMatched[2 * i], prevMatched[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher m(nnratio, checkOri != 0);
    const int n = m.SearchForInitialization(*f1, *(Frame*)f2p, prev, m12, windowSize);
    for (int i = 0; i < f1->N; i++) { prevMatched[2 * i] = prev[i].x; prevMatched[2 * i + 1] = prev[i].y; matches12[i] = m12[i]; }
    return n;
}
// ORBmatcher::SearchForTriangulation (:755-927): pairs out (2 ints each), returns the pair count (= the function's return value)
int rm_search_for_triangulation(void* kf1, void* kf2, const float* F12, int onlyStereo, int32_t* pairs, int cap, float nnratio, int checkOri)
{
    std::vector<std::pair<size_t, size_t> > vp;
    ORBmatcher m(nnratio, checkOri != 0);
    const int n = m.SearchForTriangulation((KeyFrame*)kf1, (KeyFrame*)kf2, mat32f(3, 3, F12), vp, onlyStereo != 0);
    for (size_t i = 0; i < vp.size() && (int)i < cap; i++) { pairs[2 * i] = (int32_t)vp[i].first; pairs[2 * i + 1] = (int32_t)vp[i].second; }
    return n;
}
// ORBmatcher::SearchBySim3 (:1253-1500): matches12 (in/out) = MapPointSet index of KF2's point per KF1 feature, -1 = NULL
int rm_search_by_sim3(void* kf1p, void* kf2, void* sp, int32_t* matches12, float s12, const float* R12, const float* t12, float th, float nnratio)
{
    KeyFrame* kf1 = (KeyFrame*)kf1p; MapPointSet* s = (MapPointSet*)sp;
    std::vector<MapPoint*> vm(kf1->N, (MapPoint*)nullptr);
    for (int i = 0; i < kf1->N; i++) if (matches12[i] >= 0) vm[i] = s->v[matches12[i]];
    ORBmatcher m(nnratio, true);
    const int n = m.SearchBySim3(kf1, (KeyFrame*)kf2, vm, s12, mat32f(3, 3, R12), mat32f(3, 1, t12), th);
    for (int i = 0; i < kf1->N; i++) matches12[i] = s->indexOf(vm[i]);
    return n;
}
// ORBmatcher::Fuse(KeyFrame*, const vector<MapPoint*>&, th) (:962-1115)
int rm_fuse(void* kf, void* sp, float th, float nnratio)
{
    ORBmatcher m(nnratio, true);
    return m.Fuse((KeyFrame*)kf, ((MapPointSet*)sp)->v, th);
}
// ORBmatcher::Fuse(KeyFrame*, cv::Mat Scw, const vector<MapPoint*>&, th, vector<MapPoint*>& vpReplacePoint) (:1119-1251):
// replace[i] = MapPointSet index (in `kfpoints`) of the KeyFrame's point that vpPoints[i] should replace, -1 = NULL
int rm_fuse_sim3(void* kf, const float* Scw16, void* sp, void* kfpoints, float th, int32_t* replace, float nnratio)
{
    MapPointSet* s = (MapPointSet*)sp;
    std::vector<MapPoint*> vr(s->v.size(), (MapPoint*)nullptr);
    ORBmatcher m(nnratio, true);
    const int n = m.Fuse((KeyFrame*)kf, mat32f(4, 4, Scw16), s->v, th, vr);
    for (size_t i = 0; i < vr.size(); i++) replace[i] = ((MapPointSet*)kfpoints)->indexOf(vr[i]);
    return n;
}
// ORBmatcher::DescriptorDistance (:1838-1854) for n pairs
void rm_descriptor_distance(const uint8_t* a, const uint8_t* b, int n, int32_t* out)
{
    for (int i = 0; i < n; i++) {
        cv::Mat ma(1, 32, CV_8U, (void*)(a + (size_t)i * 32)), mb(1, 32, CV_8U, (void*)(b + (size_t)i * 32));
        out[i] = ORBmatcher::DescriptorDistance(ma, mb);
    }
}

}  // extern "C"
