// oracle/cv_prims.hpp — TEST INFRASTRUCTURE (CPU oracle), never linked into the product.
//
// Plain-C++ restatement of the OpenCV primitives the reference hot path calls.  The
// arithmetic lives in a third-party dependency that is NOT vendored in /root/reference:
// OpenCV, required ">= 2.4.3" (ORB-SLAM2/CMakeLists.txt:20, README.md:60) i.e. unpinned by
// the reference.  We pin to OpenCV 4.13.0 semantics (importable here as cv2) and the tests
// in tests/test_oracle_cv2.py assert bit equality of every function below against cv2.
//
// Reference call sites (ORB-SLAM2/src/ORBextractor.cpp):
//   cv::resize INTER_LINEAR      :1136      -> resize_linear_u8
//   cv::GaussianBlur 7x7 s=2     :1097-1098 -> gaussian_blur7_u8
//   cv::FAST(cell, th, true)     :818-825   -> fast9_16_nms
//   cv::fastAtan2                :105       -> fast_atan2
//   cvRound                      :83,117-122,1128 -> cv_round
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>

namespace cvp {

// cvRound: round-half-to-even (SSE cvtss2si / cvtsd2si under the default rounding mode).
static inline int cv_round(double v) { return (int)std::nearbyint(v); }
static inline int cv_round(float v) { return (int)std::nearbyintf(v); }
static inline int cv_floor(double v) { int i = (int)v; return i - (i > v); }
static inline int cv_ceil(double v) { int i = (int)v; return i + (i < v); }

// ---------------------------------------------------------------------------------------
// cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR) for CV_8UC1 (OpenCV imgproc/resize.cpp:
// 11-bit fixed-point coefficients, HResizeLinear -> int32 rows, VResizeLinear<uchar,int,short>).
// ---------------------------------------------------------------------------------------
struct LinCoef { int s0, s1; short a0, a1; };

static inline void linear_coeffs(int src, int dst, std::vector<LinCoef>& out)
{
    out.resize(dst);
    const double inv_scale = (double)dst / src;
    const double scale = 1.0 / inv_scale;
    for (int d = 0; d < dst; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cv_floor(f);
        f -= s;
        if (s < 0) { s = 0; f = 0.f; }
        if (s >= src - 1) { s = src - 1; f = 0.f; }
        LinCoef c;
        c.s0 = s;
        c.s1 = std::min(s + 1, src - 1);
        c.a0 = (short)cv_round((1.f - f) * 2048.f);
        c.a1 = (short)cv_round(f * 2048.f);
        out[d] = c;
    }
}

static inline void resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstride,
                                    uint8_t* dst, int dw, int dh, size_t dstride)
{
    std::vector<LinCoef> cx, cy;
    linear_coeffs(sw, dw, cx);
    linear_coeffs(sh, dh, cy);
    std::vector<int> r0(dw), r1(dw);
    for (int y = 0; y < dh; y++) {
        const uint8_t* S0 = src + (size_t)cy[y].s0 * sstride;
        const uint8_t* S1 = src + (size_t)cy[y].s1 * sstride;
        for (int x = 0; x < dw; x++) {
            r0[x] = S0[cx[x].s0] * cx[x].a0 + S0[cx[x].s1] * cx[x].a1;
            r1[x] = S1[cx[x].s0] * cx[x].a0 + S1[cx[x].s1] * cx[x].a1;
        }
        const int b0 = cy[y].a0, b1 = cy[y].a1;
        uint8_t* D = dst + (size_t)y * dstride;
        for (int x = 0; x < dw; x++) {
            int v = (((b0 * (r0[x] >> 4)) >> 16) + ((b1 * (r1[x] >> 4)) >> 16) + 2) >> 2;
            D[x] = (uint8_t)std::min(std::max(v, 0), 255);
        }
    }
}

// ---------------------------------------------------------------------------------------
// cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) for CV_8UC1.
// OpenCV >= 3.4 fixed-point path (smooth.dispatch.cpp / ufixedpoint16): Q8 kernel
// {18,34,48,56,48,34,18}, horizontal pass Q8.8 (u16), vertical pass Q16.16 (u32),
// result (v + 2^15) >> 16.  legacy=true selects the OpenCV <= 3.3 float-derived integer
// kernel {18,34,49,55,49,34,18}/2^8 with the same rounding and saturation (optional mode).
// ---------------------------------------------------------------------------------------
static inline int reflect101(int p, int n)
{
    if (n == 1) return 0;
    while (p < 0 || p >= n) {
        if (p < 0) p = -p;
        else p = 2 * (n - 1) - p;
    }
    return p;
}

static inline void gaussian_blur7_u8(const uint8_t* src, int w, int h, size_t sstride,
                                     uint8_t* dst, size_t dstride, bool legacy = false)
{
    static const int Q[7] = {18, 34, 48, 56, 48, 34, 18};
    static const int QL[7] = {18, 34, 49, 55, 49, 34, 18};
    const int* q = legacy ? QL : Q;
    std::vector<uint32_t> hbuf((size_t)w * h);
    for (int y = 0; y < h; y++) {
        const uint8_t* S = src + (size_t)y * sstride;
        for (int x = 0; x < w; x++) {
            uint32_t acc = 0;
            for (int k = 0; k < 7; k++) acc += (uint32_t)q[k] * S[reflect101(x + k - 3, w)];
            hbuf[(size_t)y * w + x] = acc;
        }
    }
    for (int y = 0; y < h; y++) {
        uint8_t* D = dst + (size_t)y * dstride;
        for (int x = 0; x < w; x++) {
            uint32_t acc = 0;
            for (int k = 0; k < 7; k++) acc += (uint32_t)q[k] * hbuf[(size_t)reflect101(y + k - 3, h) * w + x];
            uint32_t v = (acc + 32768u) >> 16;
            D[x] = (uint8_t)std::min<uint32_t>(v, 255u);
        }
    }
}

// ---------------------------------------------------------------------------------------
// cv::FAST(img, kps, threshold, nonmaxSuppression=true)  == FAST-9-16 (features2d/fast.cpp,
// fast_score.cpp).  Output order is row-major; response = cornerScore = m - 1 where
// m = max over the 16 arcs of 9 contiguous ring pixels of max(min(v-p), min(p-v)).
// ---------------------------------------------------------------------------------------
struct FastKp { int x, y, score; };

static const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

// m(x,y): the threshold-independent arc score (0 if no arc has a positive value).
static inline int fast_arc_score(const uint8_t* img, size_t stride, int x, int y)
{
    const int v = img[(size_t)y * stride + x];
    int d[25];
    for (int k = 0; k < 16; k++)
        d[k] = v - img[(size_t)(y + kRingDy[k]) * stride + (x + kRingDx[k])];
    for (int k = 16; k < 25; k++) d[k] = d[k - 16];
    int best = -256;
    for (int k = 0; k < 16; k++) {
        int mn = d[k], mx = d[k];
        for (int j = 1; j < 9; j++) { mn = std::min(mn, d[k + j]); mx = std::max(mx, d[k + j]); }
        best = std::max(best, std::max(mn, -mx));
    }
    return best;
}

static inline void fast9_16_nms(const uint8_t* img, int w, int h, size_t stride, int threshold,
                                std::vector<FastKp>& out)
{
    out.clear();
    if (w < 7 || h < 7) return;
    std::vector<int> score((size_t)w * h, 0);
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            int m = fast_arc_score(img, stride, x, y);
            if (m > threshold) score[(size_t)y * w + x] = m - 1;
        }
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            int s = score[(size_t)y * w + x];
            if (s == 0) continue;   // m > th >= 0  =>  m-1 >= th; th==0 corner with m==1 scores 0 and
                                    // can never beat its (>=0) neighbours under strict '>' either.
            bool keep = true;
            for (int dy = -1; dy <= 1 && keep; dy++)
                for (int dx = -1; dx <= 1; dx++) {
                    if (!dx && !dy) continue;
                    if (s <= score[(size_t)(y + dy) * w + (x + dx)]) { keep = false; break; }
                }
            if (keep) out.push_back(FastKp{x, y, s});
        }
}

// ---------------------------------------------------------------------------------------
// cv::fastAtan2(y, x) in degrees (core/mathfuncs_core: atan_f32 polynomial, scalar path).
// Every operation is a separately rounded float op (compile with -ffp-contract=off).
// ---------------------------------------------------------------------------------------
static inline float fast_atan2(float y, float x)
{
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale;
    const float p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale;
    const float p7 = -0.04432655554792128f * scale;
    const float eps = 2.2204460492503131e-16f;   // (float)DBL_EPSILON
    float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + eps);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + eps);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

}  // namespace cvp
