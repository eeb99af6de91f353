// host_selftest.cc — exercises the C++ host mirror (ORB_SLAM2::ORBextractor / ORBmatcher) exactly as Frame.cc does.
//   host_selftest                       no arguments: API/table checks; with no CUDA device it must fail loudly (exit 0)
//   host_selftest W H in.raw out.bin    extract in.raw (W x H, 8-bit) and write n, keypoints, descriptors, level sizes
//   host_selftest match in.bin out.bin  run the ORBmatcher mirror (SearchByProjection x2, SearchForInitialization, ComputeStereoMatches,
//                                       DescriptorDistances) on dumped inputs; tests/test_host_mirror.py compares with the oracle
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdexcept>
#include "ORBextractor.h"
#include "ORBmatcher.h"

using namespace ORB_SLAM2;

static int api_checks()
{
    ORBextractor ex(1000, 1.2f, 8, 20, 7);                       // TUM1.yaml ORB settings (Tracking.cc:141-149)
    if (ex.GetLevels() != 8 || ex.GetScaleFactor() != 1.2f) return 1;
    std::vector<float> sc = ex.GetScaleFactors(), isc = ex.GetInverseScaleFactors();
    std::vector<float> s2 = ex.GetScaleSigmaSquares(), is2 = ex.GetInverseScaleSigmaSquares();
    if (sc.size() != 8 || sc[0] != 1.0f || sc[1] != 1.2f || isc[1] != 1.0f / 1.2f || s2[1] != 1.2f * 1.2f || is2[2] != 1.0f / s2[2]) return 2;
    if (ORBmatcher::TH_LOW != 50 || ORBmatcher::TH_HIGH != 100 || ORBmatcher::HISTO_LENGTH != 30) return 3;
    std::vector<cv::KeyPoint> kps;
    cv::Mat desc, empty;
    ex(empty, cv::Mat(), kps, desc);                             // empty image: silent return
    if (!kps.empty()) return 4;
    cv::Mat img(480, 640, CV_8U);
    for (int y = 0; y < 480; y++) for (int x = 0; x < 640; x++) img.ptr(y)[x] = (unsigned char)((x * 7 + y * 13) ^ (x * y >> 3));
    if (orbb200_device_count() < 1) {
        try { ex(img, cv::Mat(), kps, desc); }
        catch (const OrbDeviceError& e) {
            if (e.code != ORBB_E_CUDA) return 5;
            printf("no CUDA device: operator() failed loudly (%s)\n", e.what());
            return 0;
        }
        return 6;                                                // a silent CPU result would be a bug
    }
    ex(img, cv::Mat(), kps, desc);
    if (kps.empty() || desc.rows != (int)kps.size() || desc.cols != 32) return 7;
    if ((int)ex.mvImagePyramid.size() != 8 || ex.mvImagePyramid[1].cols != 533 || ex.mvImagePyramid[7].rows != 134) return 8;
    if (memcmp(ex.mvImagePyramid[0].ptr(5), img.ptr(5), 640)) return 9;
    int32_t d = -1;
    ORBmatcher::DescriptorDistances(desc.ptr(0), desc.ptr(0), 1, &d);
    if (d != 0) return 10;
    printf("device path ok: %zu keypoints\n", kps.size());
    return 0;
}

namespace {
struct Reader {
    FILE* f;
    template <typename T> void get(T* p, size_t n) { if (n && fread(p, sizeof(T), n, f) != n) throw std::runtime_error("short input file"); }
    int i32() { int v; get(&v, 1); return v; }
    float f32() { float v; get(&v, 1); return v; }
};
cv::Mat rows32(const std::vector<uint8_t>& d, int n)
{
    cv::Mat m(n > 0 ? n : 1, 32, CV_8U);
    for (int i = 0; i < n; i++) memcpy(m.ptr(i), &d[(size_t)i * 32], 32);
    m.rows = n;
    return m;
}
}  // namespace

// The matcher half of the mirror, driven like Tracking / Initializer / Frame drive the reference classes.
static int match_mode(const char* in, const char* outp)
{
    Reader r{fopen(in, "rb")};
    if (!r.f) { fprintf(stderr, "cannot read %s\n", in); return 2; }
    const int w = r.i32(), h = r.i32(), n1 = r.i32(), n2 = r.i32(), nq = r.i32(), nf = r.i32();
    std::vector<cv::KeyPoint> k1(n1), k2(n2);
    std::vector<uint8_t> d1((size_t)n1 * 32), d2((size_t)n2 * 32);
    r.get(k1.data(), n1); r.get(d1.data(), d1.size()); r.get(k2.data(), n2); r.get(d2.data(), d2.size());
    std::vector<float> ur2(n2); r.get(ur2.data(), n2);
    std::vector<orbm_proj_query> q(nq); r.get(q.data(), nq);
    std::vector<int32_t> mp0(n2), ob0(n2); r.get(mp0.data(), n2); r.get(ob0.data(), n2);
    std::vector<cv::Point2f> prev(n1); r.get(prev.data(), n1);
    cv::Mat imL(h, w, CV_8U), imR(h, w, CV_8U);
    r.get(imL.ptr(0), (size_t)w * h); r.get(imR.ptr(0), (size_t)w * h);
    const float mbf = r.f32();
    fclose(r.f);
    FILE* o = fopen(outp, "wb");
    auto put = [&](const void* p, size_t bytes) { fwrite(p, 1, bytes, o); };

    const cv::Mat D1 = rows32(d1, n1), D2 = rows32(d2, n2);
    DeviceFrame F1(k1, D1, 0.f, 0.f, (float)w, (float)h), F2(k2, D2, 0.f, 0.f, (float)w, (float)h);
    {   // Tracking::TrackWithMotionModel -> SearchByProjection(CurrentFrame, LastFrame, th, mono) (src/ORBmatcher.cpp:1506)
        std::vector<int32_t> mp = mp0, ob = ob0;
        const int n = ORBmatcher(0.9f, true).SearchByProjectionLastFrame(F2, &ur2, q, mp, ob);
        put(&n, 4); put(mp.data(), 4 * (size_t)n2); put(ob.data(), 4 * (size_t)n2);
    }
    {   // Tracking::SearchLocalPoints -> SearchByProjection(CurrentFrame, vpMapPoints, th) (:63)
        std::vector<int32_t> mp = mp0, ob = ob0;
        const int n = ORBmatcher(0.8f).SearchByProjection(F2, &ur2, q, mp, ob);
        put(&n, 4); put(mp.data(), 4 * (size_t)n2); put(ob.data(), 4 * (size_t)n2);
    }
    {   // Tracking::MonocularInitialization -> SearchForInitialization(mInitialFrame, mCurrentFrame, mvbPrevMatched, mvIniMatches, 100) (:474)
        std::vector<int> m12;
        std::vector<cv::Point2f> pm = prev;
        const int n = ORBmatcher(0.9f, true).SearchForInitialization(F1, F2, pm, m12, 100);
        put(&n, 4); put(m12.data(), 4 * (size_t)n1); put(pm.data(), 8 * (size_t)n1);
    }
    {   // Frame::Frame (stereo): two extractors, then ComputeStereoMatches (src/Frame.cpp:84-100, :540)
        ORBextractor exL(nf, 1.2f, 8, 20, 7), exR(nf, 1.2f, 8, 20, 7);
        exL.SetPyramidDownload(false); exR.SetPyramidDownload(false);
        std::vector<cv::KeyPoint> kL, kR;
        cv::Mat dL, dR;
        exL(imL, cv::Mat(), kL, dL); exR(imR, cv::Mat(), kR, dR);
        std::vector<float> uR, depth;
        ComputeStereoMatches(exL, exR, kL, dL, kR, dR, mbf, 0.0f, uR, depth);
        const int n = (int)kL.size();
        put(&n, 4); put(uR.data(), 4 * (size_t)n); put(depth.data(), 4 * (size_t)n);
        // the same constructor as ONE device call (orbm_stereo_frame_create): every member must come back identical, and a tracked-frame
        // search on the frame it leaves on the device must not depend on where mvuRight comes from
        std::vector<cv::KeyPoint> kL2, kR2;
        cv::Mat dL2, dR2;
        std::vector<float> uR2, depth2;
        DeviceFrame F(exL, imL, imR, mbf, 0.0f, 0.f, 0.f, (float)w, (float)h, kL2, dL2, kR2, dR2, uR2, depth2);
        if (kL2.size() != kL.size() || kR2.size() != kR.size() || memcmp(kL2.data(), kL.data(), sizeof(cv::KeyPoint) * kL.size()) ||
            memcmp(kR2.data(), kR.data(), sizeof(cv::KeyPoint) * kR.size()) || memcmp(uR2.data(), uR.data(), 4 * uR.size()) ||
            memcmp(depth2.data(), depth.data(), 4 * depth.size())) { fprintf(stderr, "fused stereo constructor differs (keypoints / uRight / depth)\n"); return 4; }
        for (int i = 0; i < n; i++) if (memcmp(dL2.ptr(i), dL.ptr(i), 32)) { fprintf(stderr, "fused stereo constructor differs (descriptor %d)\n", i); return 4; }
        std::vector<orbm_proj_query> qs((size_t)n);
        for (int i = 0; i < n; i++) {
            orbm_proj_query& e = qs[i];
            memset(&e, 0, sizeof e);
            e.q.u = kL[i].pt.x + 2.0f; e.q.v = kL[i].pt.y + 1.0f; e.q.r = 15.0f * exL.GetScaleFactors()[kL[i].octave];
            e.q.min_level = kL[i].octave - 1; e.q.max_level = kL[i].octave + 1; memcpy(e.q.desc, dL.ptr(i), 32);
            e.u_right = e.q.u - 5.0f; e.max_err_right = e.q.r; e.id = i; e.has_obs = 1; e.angle = kL[i].angle;
        }
        std::vector<int32_t> mpA(n, -1), obA(n, 0), mpB(n, -1), obB(n, 0);
        const int nA = ORBmatcher(0.9f, true).SearchByProjectionLastFrame(F, DeviceFrame::StoredURight(), qs, mpA, obA);
        const int nB = ORBmatcher(0.9f, true).SearchByProjectionLastFrame(F, &uR, qs, mpB, obB);
        if (nA != nB || nA < n / 4 || mpA != mpB || obA != obB) { fprintf(stderr, "search on the fused frame: %d vs %d matches\n", nA, nB); return 4; }
    }
    {   // Frame::Frame (monocular) as one device call against ExtractORB + UndistortKeyPoints done member by member (TUM1.yaml calibration)
        const float K4[4] = {517.306408f, 516.469215f, 318.643040f, 255.313989f};
        const std::vector<float> D = {0.262383f, -0.953104f, -0.005358f, 0.002628f, 1.163314f};
        ORBextractor ex(nf, 1.2f, 8, 20, 7);
        std::vector<cv::KeyPoint> k, ku, k2, ku2;
        cv::Mat d, d2;
        ex(imL, cv::Mat(), k, d);
        UndistortKeyPoints(k, K4, D, ku);
        float bx0, by0, bx1, by1;
        ComputeImageBounds(w, h, K4, D, bx0, by0, bx1, by1);
        DeviceFrame F(ex, imL, K4, D, bx0, by0, bx1, by1, k2, ku2, d2);
        if (k2.size() != k.size() || memcmp(k2.data(), k.data(), sizeof(cv::KeyPoint) * k.size()) || memcmp(ku2.data(), ku.data(), sizeof(cv::KeyPoint) * k.size()))
        { fprintf(stderr, "fused monocular constructor differs (keypoints)\n"); return 5; }
        for (size_t i = 0; i < k.size(); i++) if (memcmp(d2.ptr((int)i), d.ptr((int)i), 32)) { fprintf(stderr, "fused monocular constructor differs (descriptor %zu)\n", i); return 5; }
        if (F.size() != (int)k.size()) return 5;
    }
    {   // batched distances: row i of frame 1 against row i of frame 2
        const int n = n1 < n2 ? n1 : n2;
        std::vector<int32_t> dist(n);
        ORBmatcher::DescriptorDistances(d1.data(), d2.data(), n, dist.data());
        put(&n, 4); put(dist.data(), 4 * (size_t)n);
    }
    fclose(o);
    return 0;
}

int main(int argc, char** argv)
{
    if (argc == 4 && !strcmp(argv[1], "match")) {
        try { return match_mode(argv[2], argv[3]); }
        catch (const std::exception& e) { fprintf(stderr, "host_selftest match: %s\n", e.what()); return 3; }
    }
    if (argc < 5) { const int rc = api_checks(); if (rc) fprintf(stderr, "host_selftest: check %d failed\n", rc); return rc; }
    const int W = atoi(argv[1]), H = atoi(argv[2]);
    const int nf = argc > 5 ? atoi(argv[5]) : 1000;
    cv::Mat img(H, W, CV_8U);
    FILE* f = fopen(argv[3], "rb");
    if (!f || fread(img.ptr(0), 1, (size_t)W * H, f) != (size_t)W * H) { fprintf(stderr, "cannot read %s\n", argv[3]); return 2; }
    fclose(f);
    try {
        ORBextractor ex(nf, 1.2f, 8, 20, 7);
        std::vector<cv::KeyPoint> kps;
        cv::Mat desc;
        ex(img, cv::Mat(), kps, desc);
        FILE* o = fopen(argv[4], "wb");
        int n = (int)kps.size();
        fwrite(&n, 4, 1, o);
        fwrite(kps.data(), sizeof(cv::KeyPoint), n, o);
        for (int i = 0; i < n; i++) fwrite(desc.ptr(i), 1, 32, o);
        for (int l = 0; l < 8; l++) {
            int wh[2] = {ex.mvImagePyramid[l].cols, ex.mvImagePyramid[l].rows};
            fwrite(wh, 4, 2, o);
            for (int y = 0; y < wh[1]; y++) fwrite(ex.mvImagePyramid[l].ptr(y), 1, wh[0], o);
        }
        fclose(o);
    } catch (const std::exception& e) { fprintf(stderr, "host_selftest: %s\n", e.what()); return 3; }
    return 0;
}
