// oracle/oracle_capi.cpp — TEST INFRASTRUCTURE.  C API (ctypes) over the CPU oracle.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference legs may load this.
#include "orb_extract_oracle.hpp"
#include "orb_match_oracle.hpp"
#include "dbow_oracle.hpp"
#include <chrono>
#include <thread>

using namespace orbo;

extern "C" {

struct orc_ctx {
    Extractor ex;
    std::vector<KeyPoint> kps;
    std::vector<uint8_t> desc;
    orc_ctx(int nf, float sf, int nl, int ini, int mn) : ex(nf, sf, nl, ini, mn) {}
};

orc_ctx* orc_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, int tieMode, int legacyBlur)
{
    orc_ctx* c = new orc_ctx(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    c->ex.tieMode = tieMode; c->ex.legacyBlur = legacyBlur != 0;
    return c;
}
void orc_destroy(orc_ctx* c) { delete c; }

int orc_extract(orc_ctx* c, const uint8_t* img, int w, int h, long stride)
{
    c->ex.run(img, w, h, (size_t)stride, c->kps, c->desc);
    return (int)c->kps.size();
}
int orc_get_result(orc_ctx* c, KeyPoint* kps, uint8_t* desc, int cap)
{
    int n = std::min<int>(cap, (int)c->kps.size());
    if (n) { memcpy(kps, c->kps.data(), sizeof(KeyPoint) * n); memcpy(desc, c->desc.data(), (size_t)32 * n); }
    return (int)c->kps.size();
}
void orc_tables(orc_ctx* c, float* scale, float* invScale, float* sigma2, float* invSigma2, int* perLevel, int* umax)
{
    for (int i = 0; i < c->ex.nlevels; i++) {
        scale[i] = c->ex.mvScaleFactor[i]; invScale[i] = c->ex.mvInvScaleFactor[i];
        sigma2[i] = c->ex.mvLevelSigma2[i]; invSigma2[i] = c->ex.mvInvLevelSigma2[i];
        perLevel[i] = c->ex.mnFeaturesPerLevel[i];
    }
    for (int i = 0; i < 16; i++) umax[i] = c->ex.umax[i];
}
void orc_level_size(orc_ctx* c, int level, int* w, int* h) { *w = c->ex.pyramid[level].w; *h = c->ex.pyramid[level].h; }
void orc_level_image(orc_ctx* c, int level, int blurred, uint8_t* dst)
{
    const Image& im = blurred ? c->ex.blurred[level] : c->ex.pyramid[level];
    memcpy(dst, im.px.data(), im.px.size());
}
int orc_level_raw_count(orc_ctx* c, int level) { return (int)c->ex.dbg[level].toDistribute.size(); }
void orc_level_raw(orc_ctx* c, int level, float* xyr)
{
    const auto& v = c->ex.dbg[level].toDistribute;
    for (size_t i = 0; i < v.size(); i++) { xyr[3 * i] = v[i].x; xyr[3 * i + 1] = v[i].y; xyr[3 * i + 2] = v[i].response; }
}
int orc_level_kp_count(orc_ctx* c, int level) { return (int)c->ex.dbg[level].kps.size(); }
void orc_level_kps(orc_ctx* c, int level, KeyPoint* out)
{
    const auto& v = c->ex.dbg[level].kps;
    if (!v.empty()) memcpy(out, v.data(), v.size() * sizeof(KeyPoint));
}

// ---- primitives (for the cv2 pinning tests) ---------------------------------------------
void orc_resize(const uint8_t* src, int sw, int sh, uint8_t* dst, int dw, int dh) { cvp::resize_linear_u8(src, sw, sh, sw, dst, dw, dh, dw); }
void orc_remap(const uint8_t* src, int sw, int sh, const float* mapx, const float* mapy, uint8_t* dst, int dw, int dh) { cvp::remap_linear_u8(src, sw, sh, sw, mapx, mapy, dst, dw, dh, dw); }
void orc_cvt_gray(const uint8_t* src, int w, int h, int channels, int rgb_order, uint8_t* dst) { cvp::cvt_gray_u8(src, w, h, (size_t)w * channels, channels, rgb_order != 0, dst, w); }
void orc_undistort_points(const float* src_xy, int n, const float* K4, const float* dist, int nd, float* dst_xy) { cvp::undistort_points_f32(src_xy, n, K4[0], K4[1], K4[2], K4[3], dist, nd, dst_xy); }
// Frame::UndistortKeyPoints (Frame.cpp:462-496): copy when mDistCoef(0) == 0, else cv::undistortPoints on the (x, y) pairs.
void orc_undistort_keypoints(const KeyPoint* kps, int n, const float* K4, const float* dist, int nd, KeyPoint* out)
{
    for (int i = 0; i < n; i++) out[i] = kps[i];
    if (dist[0] == 0.0) return;
    std::vector<float> xy(2 * (size_t)n + 2), un(2 * (size_t)n + 2);
    for (int i = 0; i < n; i++) { xy[2 * i] = kps[i].x; xy[2 * i + 1] = kps[i].y; }
    cvp::undistort_points_f32(xy.data(), n, K4[0], K4[1], K4[2], K4[3], dist, nd, un.data());
    for (int i = 0; i < n; i++) { out[i].x = un[2 * i]; out[i].y = un[2 * i + 1]; }
}
// Frame::ComputeImageBounds (Frame.cpp:498-536) -> {mnMinX, mnMinY, mnMaxX, mnMaxY}
void orc_image_bounds(int cols, int rows, const float* K4, const float* dist, int nd, float* b)
{
    if (dist[0] != 0.0) {
        const float c[8] = {0.f, 0.f, (float)cols, 0.f, 0.f, (float)rows, (float)cols, (float)rows};
        float u[8];
        cvp::undistort_points_f32(c, 4, K4[0], K4[1], K4[2], K4[3], dist, nd, u);
        b[0] = std::min(u[0], u[4]); b[2] = std::max(u[2], u[6]); b[1] = std::min(u[1], u[3]); b[3] = std::max(u[5], u[7]);
    } else { b[0] = 0.0f; b[2] = (float)cols; b[1] = 0.0f; b[3] = (float)rows; }
}
void orc_blur(const uint8_t* src, int w, int h, uint8_t* dst, int legacy) { cvp::gaussian_blur7_u8(src, w, h, w, dst, w, legacy != 0); }
int orc_fast(const uint8_t* img, int w, int h, int th, int* xys, int cap)
{
    std::vector<cvp::FastKp> v;
    cvp::fast9_16_nms(img, w, h, w, th, v);
    for (int i = 0; i < (int)v.size() && i < cap; i++) { xys[3 * i] = v[i].x; xys[3 * i + 1] = v[i].y; xys[3 * i + 2] = v[i].score; }
    return (int)v.size();
}
void orc_fast_atan2(const float* y, const float* x, float* out, int n) { for (int i = 0; i < n; i++) out[i] = cvp::fast_atan2(y[i], x[i]); }

// ---- matching -----------------------------------------------------------------------------
void orc_distance_pairs(const uint8_t* a, const uint8_t* b, int n, int* out)
{
    for (int i = 0; i < n; i++) out[i] = DescriptorDistance(a + (size_t)i * 32, b + (size_t)i * 32);
}
void orc_distinctive(const uint8_t* desc, const int64_t* off, int npoints, int32_t* best_row, int32_t* best_median)
{
    for (int p = 0; p < npoints; p++) {
        int med = -1;
        best_row[p] = DistinctiveDescriptor(desc + (size_t)off[p] * 32, (int)(off[p + 1] - off[p]), &med);
        best_median[p] = med;
    }
}
void orc_bf_top2(const uint8_t* db, long M, const uint8_t* q, int nq, int32_t* d1, int64_t* i1, int32_t* d2, int nthreads)
{
    if (nthreads <= 1) { bf_top2(db, M, q, nq, d1, i1, d2); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++) {
        int q0 = (int)((long)nq * t / nthreads), q1 = (int)((long)nq * (t + 1) / nthreads);
        th.emplace_back([=]() { bf_top2(db, M, q + (size_t)q0 * 32, q1 - q0, d1 + q0, i1 + q0, d2 + q0); });
    }
    for (auto& t : th) t.join();
}

// ---- DBoW2 vocabulary descent (Frame::ComputeBoW) ---------------------------------------------
Vocabulary* orc_vocab_create(int k, int L, int n, const int32_t* parent, const uint8_t* is_leaf, const uint8_t* desc, const double* weight,
                             int weighting, int norm)
{
    Vocabulary* v = new Vocabulary();
    v->build(k, L, n, parent, is_leaf, desc, weight);
    v->weighting = weighting; v->norm = norm;
    return v;
}
void orc_vocab_destroy(Vocabulary* v) { delete v; }
void orc_vocab_transform(const Vocabulary* v, const uint8_t* feats, int n, int levelsup, int32_t* word_id, double* weight, int32_t* node_id)
{
    for (int i = 0; i < n; i++) {
        int id, nid; double w;
        v->transform(feats + (size_t)i * 32, id, w, &nid, levelsup);
        word_id[i] = id; weight[i] = w; node_id[i] = nid;
    }
}
// BowVector as (ids, values) ascending; FeatureVector as CSR (node ids ascending, offsets, feature indices in insertion order)
void orc_vocab_bow(const Vocabulary* v, const uint8_t* feats, int n, int levelsup, int32_t* bow_ids, double* bow_vals, int32_t* n_bow,
                   int32_t* fv_nodes, int32_t* fv_off, int32_t* fv_feat, int32_t* n_fv)
{
    std::map<unsigned, double> bv;
    std::map<unsigned, std::vector<unsigned> > fv;
    v->transform_all(feats, n, levelsup, bv, fv);
    int i = 0;
    for (auto& e : bv) { bow_ids[i] = (int32_t)e.first; bow_vals[i] = e.second; i++; }
    *n_bow = i;
    int j = 0, pos = 0;
    fv_off[0] = 0;
    for (auto& e : fv) {
        fv_nodes[j] = (int32_t)e.first;
        for (unsigned f : e.second) fv_feat[pos++] = (int32_t)f;
        fv_off[++j] = pos;
    }
    *n_fv = j;
}

static std::map<unsigned, std::vector<unsigned> > fv_from_csr(const int32_t* nodes, const int32_t* off, const int32_t* feat, int n)
{
    std::map<unsigned, std::vector<unsigned> > fv;
    for (int j = 0; j < n; j++) fv[(unsigned)nodes[j]] = std::vector<unsigned>(feat + off[j], feat + off[j + 1]);
    return fv;
}
int orc_search_by_bow(const uint8_t* descKF, const float* angleKF, const uint8_t* kf_has_mp,
                      const int32_t* kn, const int32_t* ko, const int32_t* kf, int nk,
                      const uint8_t* descF, const float* angleF, int FN, const int32_t* fn, const int32_t* fo, const int32_t* ff, int nf,
                      float nnratio, int check_ori, int32_t* matches)
{
    std::vector<int> m;
    const int n = SearchByBoW(descKF, angleKF, kf_has_mp, fv_from_csr(kn, ko, kf, nk), descF, angleF, FN, fv_from_csr(fn, fo, ff, nf),
                              nnratio, check_ori != 0, m);
    for (int i = 0; i < FN; i++) matches[i] = m[i];
    return n;
}

int orc_search_by_bow_kf(const uint8_t* d1, const float* a1, const uint8_t* h1, int n1, const int32_t* n1n, const int32_t* n1o, const int32_t* n1f, int nf1,
                         const uint8_t* d2, const float* a2, const uint8_t* h2, int n2, const int32_t* n2n, const int32_t* n2o, const int32_t* n2f, int nf2,
                         float nnratio, int check_ori, int32_t* matches)
{
    std::vector<int> m;
    const int n = SearchByBoW_KF(d1, a1, h1, n1, fv_from_csr(n1n, n1o, n1f, nf1), d2, a2, h2, n2, fv_from_csr(n2n, n2o, n2f, nf2), nnratio, check_ori != 0, m);
    for (int i = 0; i < n1; i++) matches[i] = m[i];
    return n;
}

int orc_search_for_triangulation(const uint8_t* d1, const KeyPoint* k1, const float* ur1, const uint8_t* h1, int n1, const int32_t* a0, const int32_t* a1, const int32_t* a2, int na,
                                 const uint8_t* d2, const KeyPoint* k2, const float* ur2, const uint8_t* h2, int n2, const int32_t* b0, const int32_t* b1, const int32_t* b2, int nb,
                                 const float* F12, float ex, float ey, const float* sf2, const float* ls2, int onlyStereo, int checkOri, int32_t* matches)
{
    std::vector<int> m;
    const int n = SearchForTriangulation(d1, k1, ur1, h1, n1, fv_from_csr(a0, a1, a2, na), d2, k2, ur2, h2, n2, fv_from_csr(b0, b1, b2, nb),
                                         F12, ex, ey, sf2, ls2, onlyStereo != 0, checkOri != 0, m);
    for (int i = 0; i < n1; i++) matches[i] = m[i];
    return n;
}

struct orc_frame { FrameGrid g; std::vector<KeyPoint> kps; std::vector<uint8_t> desc; };
orc_frame* orc_frame_create(const KeyPoint* kps, const uint8_t* desc, int n, float minx, float miny, float maxx, float maxy)
{
    orc_frame* f = new orc_frame;
    f->kps.assign(kps, kps + n); f->desc.assign(desc, desc + (size_t)n * 32);
    f->g.build(f->kps.data(), f->desc.data(), n, minx, miny, maxx, maxy);
    return f;
}
void orc_frame_destroy(orc_frame* f) { delete f; }
int orc_features_in_area(orc_frame* f, float x, float y, float r, int minLevel, int maxLevel, int* out, int cap)
{
    std::vector<int> v;
    f->g.GetFeaturesInArea(x, y, r, minLevel, maxLevel, v);
    for (int i = 0; i < (int)v.size() && i < cap; i++) out[i] = v[i];
    return (int)v.size();
}
// CSR candidate lists with distances, reference enumeration order (the contract of orbm_window_search)
long orc_window_search(orc_frame* f, const ProjQuery* qs, int nq, int* offsets, int* idx, int* dist, long cap)
{
    std::vector<int> v; long total = 0;
    for (int i = 0; i < nq; i++) {
        offsets[i] = (int)total;
        f->g.GetFeaturesInArea(qs[i].u, qs[i].v, qs[i].r, qs[i].minLevel, qs[i].maxLevel, v);
        for (int c : v) {
            if (total < cap) { idx[total] = c; dist[total] = DescriptorDistance(qs[i].desc, f->desc.data() + (size_t)c * 32); }
            total++;
        }
    }
    offsets[nq] = (int)total;
    return total;
}
int orc_search_local_points(orc_frame* f, const float* mvuRight, const ProjQuery* qs, int nq, float nnratio, int* frameMP, int* frameMPObs)
{
    std::vector<ProjQuery> q(qs, qs + nq);
    std::vector<int> mp(frameMP, frameMP + f->g.N), ob(frameMPObs, frameMPObs + f->g.N);
    int n = SearchByProjection_LocalPoints(f->g, mvuRight, q, nnratio, mp, ob);
    std::copy(mp.begin(), mp.end(), frameMP); std::copy(ob.begin(), ob.end(), frameMPObs);
    return n;
}
int orc_search_last_frame(orc_frame* f, const float* mvuRight, const ProjQuery* qs, int nq, int checkOri, int* frameMP, int* frameMPObs)
{
    std::vector<ProjQuery> q(qs, qs + nq);
    std::vector<int> mp(frameMP, frameMP + f->g.N), ob(frameMPObs, frameMPObs + f->g.N);
    int n = SearchByProjection_LastFrame(f->g, mvuRight, q, checkOri != 0, mp, ob);
    std::copy(mp.begin(), mp.end(), frameMP); std::copy(ob.begin(), ob.end(), frameMPObs);
    return n;
}
int orc_search_reloc(orc_frame* f, const ProjQuery* qs, int nq, int orbdist, int checkOri, int* frameMP)
{
    std::vector<ProjQuery> q(qs, qs + nq);
    std::vector<int> mp(frameMP, frameMP + f->g.N);
    int n = SearchByProjection_Reloc(f->g, q, orbdist, checkOri != 0, mp);
    std::copy(mp.begin(), mp.end(), frameMP);
    return n;
}
int orc_search_loop(orc_frame* f, const ProjQuery* qs, int nq, int* matched)
{
    std::vector<ProjQuery> q(qs, qs + nq);
    std::vector<int> m(matched, matched + f->g.N);
    int n = SearchByProjection_Loop(f->g, q, m);
    std::copy(m.begin(), m.end(), matched);
    return n;
}
void orc_fuse_best(orc_frame* f, const ProjQuery* qs, int nq, const float* mvuRight, const float* invSigma2, int gate, int maxDist, int* bestIdx, int* bestDist)
{
    std::vector<ProjQuery> q(qs, qs + nq);
    std::vector<int> bi, bd;
    FuseBest(f->g, q, mvuRight, invSigma2, gate != 0, maxDist, bi, bd);
    std::copy(bi.begin(), bi.end(), bestIdx); std::copy(bd.begin(), bd.end(), bestDist);
}
int orc_search_for_initialization(orc_frame* f1, orc_frame* f2, float* prevMatchedXY, int* vnMatches12, int windowSize, float nnratio, int checkOri)
{
    std::vector<float> pm(prevMatchedXY, prevMatchedXY + 2 * f1->g.N);
    std::vector<int> m12;
    int n = SearchForInitialization(f1->g, f2->g, pm, m12, windowSize, nnratio, checkOri != 0);
    std::copy(pm.begin(), pm.end(), prevMatchedXY); std::copy(m12.begin(), m12.end(), vnMatches12);
    return n;
}
// Stereo: both extractors must have been run (pyramids are read from the contexts).
void orc_stereo_match(orc_ctx* L, orc_ctx* R, float mbf, float mb, float* uRight, float* depth, int* bestIdxR, int* bestDist)
{
    ComputeStereoMatches(L->kps.data(), L->desc.data(), (int)L->kps.size(), R->kps.data(), R->desc.data(), (int)R->kps.size(),
                         L->ex.pyramid, R->ex.pyramid, L->ex.mvScaleFactor, L->ex.mvInvScaleFactor, mbf, mb, uRight, depth, bestIdxR, bestDist);
}

}  // extern "C"
