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
// and, for the SURVEY 8(f)-4 "next" row (ORB-SLAM2/src/Tracking.cpp:202-227, 233-250, 264-281):
//   cv::cvtColor RGB/BGR/RGBA/BGRA -> GRAY    -> cvt_gray_u8
//   cv::remap INTER_LINEAR (stereo rectification, Examples/Stereo/stereo_euroc.cc:174-185) -> remap_linear_u8
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <cstddef>
#include <vector>
#include <algorithm>

namespace cvp {

// cvRound: round-half-to-even (SSE cvtss2si / cvtsd2si under the default rounding mode).
static inline int cv_round(double v) { return (int)std::nearbyint(v); }
static inline int cv_round(float v) { return (int)std::nearbyintf(v); }
static inline int cv_floor(double v) { int i = (int)v; return i - (i > v); }
static inline int cv_ceil(double v) { int i = (int)v; return i + (i < v); }

// ---------------------------------------------------------------------------------------
// cv::cvtColor(src, dst, CV_RGB2GRAY | CV_BGR2GRAY | CV_RGBA2GRAY | CV_BGRA2GRAY) for 8-bit images (OpenCV 4.x
// imgproc/color_rgb: RGB2Gray<uchar>, 15-bit fixed point): gray = (R*9798 + G*19235 + B*3735 + 2^14) >> 15; alpha ignored.
// Checked against cv2 4.13 on all 2^24 colours (tests/golden/cv2_cvtcolor.npz holds the digest).
// ---------------------------------------------------------------------------------------
static inline void cvt_gray_u8(const uint8_t* src, int w, int h, size_t sstride, int channels, bool rgb_order,
                               uint8_t* dst, size_t dstride)
{
    const int RY = 9798, GY = 19235, BY = 3735, shift = 15;
    for (int y = 0; y < h; y++) {
        const uint8_t* s = src + (size_t)y * sstride;
        uint8_t* d = dst + (size_t)y * dstride;
        for (int x = 0; x < w; x++, s += channels) {
            const int r = rgb_order ? s[0] : s[2], g = s[1], b = rgb_order ? s[2] : s[0];
            d[x] = (uint8_t)((r * RY + g * GY + b * BY + (1 << (shift - 1))) >> shift);
        }
    }
}

// ---------------------------------------------------------------------------------------
// cv::remap(src, dst, map1, map2, cv::INTER_LINEAR) for CV_8UC1 with CV_32FC1 maps and the default BORDER_CONSTANT (0)
// (OpenCV imgproc/imgwarp.cpp: maps are converted to fixed point with INTER_BITS = 5, sx = cvRound(mapx * 32), integer part
// sx >> 5 saturated to short, fraction sx & 31; BilinearTab_i weights are (32-fx)(32-fy)*32 etc. — products of 1/32 steps are
// exact, so the table needs no fix-up — and the result is (sum w*p + 2^14) >> 15).  Checked against cv2 4.13
// (tests/golden/cv2_remap.npz).
// ---------------------------------------------------------------------------------------
static inline void remap_linear_u8(const uint8_t* src, int sw, int sh, size_t sstride, const float* mapx, const float* mapy,
                                   uint8_t* dst, int dw, int dh, size_t dstride)
{
    for (int y = 0; y < dh; y++)
        for (int x = 0; x < dw; x++) {
            const int sx = cv_round(mapx[(size_t)y * dw + x] * 32.f), sy = cv_round(mapy[(size_t)y * dw + x] * 32.f);
            int ix = sx >> 5, iy = sy >> 5;
            ix = ix < -32768 ? -32768 : (ix > 32767 ? 32767 : ix);
            iy = iy < -32768 ? -32768 : (iy > 32767 ? 32767 : iy);
            const int fx = sx & 31, fy = sy & 31;
            int acc = 0;
            for (int dy = 0; dy < 2; dy++)
                for (int dx = 0; dx < 2; dx++) {
                    const int xx = ix + dx, yy = iy + dy;
                    const int p = (xx >= 0 && xx < sw && yy >= 0 && yy < sh) ? src[(size_t)yy * sstride + xx] : 0;
                    const int w = (dx ? fx : 32 - fx) * (dy ? fy : 32 - fy) * 32;
                    acc += w * p;
                }
            dst[(size_t)y * dstride + x] = (uint8_t)((acc + (1 << 14)) >> 15);
        }
}

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
    std::vector<uint16_t> hbuf((size_t)w * h);
    std::vector<uint8_t> pad((size_t)w + 6);
    for (int y = 0; y < h; y++) {
        const uint8_t* S = src + (size_t)y * sstride;
        for (int x = 0; x < w + 6; x++) pad[x] = S[reflect101(x - 3, w)];
        uint16_t* H = &hbuf[(size_t)y * w];
        for (int x = 0; x < w; x++) {
            const uint8_t* p = &pad[x];
            H[x] = (uint16_t)(q[0] * p[0] + q[1] * p[1] + q[2] * p[2] + q[3] * p[3] + q[4] * p[4] + q[5] * p[5] + q[6] * p[6]);
        }
    }
    for (int y = 0; y < h; y++) {
        const uint16_t* R[7];
        for (int k = 0; k < 7; k++) R[k] = &hbuf[(size_t)reflect101(y + k - 3, h) * w];
        uint8_t* D = dst + (size_t)y * dstride;
        for (int x = 0; x < w; x++) {
            const uint32_t acc = (uint32_t)q[0] * R[0][x] + (uint32_t)q[1] * R[1][x] + (uint32_t)q[2] * R[2][x] + (uint32_t)q[3] * R[3][x] +
                                 (uint32_t)q[4] * R[4][x] + (uint32_t)q[5] * R[5][x] + (uint32_t)q[6] * R[6][x];
            const uint32_t v = (acc + 32768u) >> 16;
            D[x] = (uint8_t)(v > 255u ? 255u : v);
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
    int d[16], e[16];
    for (int k = 0; k < 16; k++) { d[k] = v - img[(size_t)(y + kRingDy[k]) * stride + (x + kRingDx[k])]; e[k] = -d[k]; }
    int a2[16], b2[16], a4[16], b4[16];
    for (int k = 0; k < 16; k++) { a2[k] = std::min(d[k], d[(k + 1) & 15]); b2[k] = std::min(e[k], e[(k + 1) & 15]); }
    for (int k = 0; k < 16; k++) { a4[k] = std::min(a2[k], a2[(k + 2) & 15]); b4[k] = std::min(b2[k], b2[(k + 2) & 15]); }
    int best = -256;
    for (int k = 0; k < 16; k++) {
        const int a9 = std::min(std::min(a4[k], a4[(k + 4) & 15]), d[(k + 8) & 15]);
        const int b9 = std::min(std::min(b4[k], b4[(k + 4) & 15]), e[(k + 8) & 15]);
        best = std::max(best, std::max(a9, b9));
    }
    return best;
}

static inline void fast9_16_nms(const uint8_t* img, int w, int h, size_t stride, int threshold,
                                std::vector<FastKp>& out)
{
    out.clear();
    if (w < 7 || h < 7) return;
    std::vector<uint8_t> score((size_t)w * h, 0);
    std::vector<int> cand;
    std::vector<uint8_t> flag((size_t)w, 0);
    const int t = threshold;
    std::ptrdiff_t off[16];
    for (int k = 0; k < 16; k++) off[k] = (std::ptrdiff_t)kRingDy[k] * (std::ptrdiff_t)stride + kRingDx[k];
    for (int y = 3; y < h - 3; y++) {
        const uint8_t* r0 = img + (size_t)y * stride;
        const uint8_t* rm = r0 - 3 * stride;
        const uint8_t* rp = r0 + 3 * stride;
        // pass 1 (auto-vectorised): necessary condition, as cv::FAST's first tests — any 9-arc holds two ADJACENT
        // compass pixels of one polarity
        for (int x = 3; x < w - 3; x++) {
            const int v = r0[x];
            const int d0 = v - rp[x], d4 = v - r0[x + 3], d8 = v - rm[x], d12 = v - r0[x - 3];
            const int k0 = d0 > t, k4 = d4 > t, k8 = d8 > t, k12 = d12 > t;
            const int b0 = -d0 > t, b4 = -d4 > t, b8 = -d8 > t, b12 = -d12 > t;
            flag[x] = (uint8_t)((k0 & k4) | (k4 & k8) | (k8 & k12) | (k12 & k0) | (b0 & b4) | (b4 & b8) | (b8 & b12) | (b12 & b0));
        }
        // pass 2: exact corner test on 16-bit brighter/darker masks, score only for corners
        for (int x = 3; x < w - 3; x++) {
            if (!flag[x]) continue;
            const uint8_t* pc = r0 + x;
            const int v = *pc, lo = v - t, hi = v + t;
            unsigned dk = 0, br = 0;
            for (int k = 0; k < 16; k++) {
                const int p = pc[off[k]];
                dk |= (unsigned)(p < lo) << k; br |= (unsigned)(p > hi) << k;
            }
            unsigned a = dk | (dk << 16), b = br | (br << 16);
            unsigned ya = a & (a >> 1), yb = b & (b >> 1);
            ya &= ya >> 2; yb &= yb >> 2; ya &= ya >> 4; yb &= yb >> 4; ya &= a >> 8; yb &= b >> 8;
            if (!((ya | yb) & 0xffffu)) continue;
            const int m = fast_arc_score(img, stride, x, y);
            score[(size_t)y * w + x] = (uint8_t)(m - 1); cand.push_back(y * w + x);
        }
    }
    for (int idx : cand) {
        const int y = idx / w, x = idx - y * w;
        const int s = score[idx];
        if (s == 0) continue;   // threshold 0 with m == 1: response 0 never beats its neighbours under strict '>'
        const uint8_t* c = &score[idx];
        if (s > c[-1] && s > c[1] && s > c[-w - 1] && s > c[-w] && s > c[-w + 1] && s > c[w - 1] && s > c[w] && s > c[w + 1])
            out.push_back(FastKp{x, y, s});
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

// cv::undistortPoints(src, dst, K, D, noArray(), P = K) for CV_32FC2 points, OpenCV 4.x (imgproc/src/undistort.dispatch.cpp,
// cvUndistortPointsInternal; the 6-argument overload uses TermCriteria(MAX_ITER, 5, 0.01), i.e. exactly five fixed-point iterations).
// All arithmetic is double on operands converted from float; R = identity and P = K make RR = K exactly; no tilt (k[12], k[13] = 0).
// dist holds nd = 4, 5 or 8 coefficients (k1 k2 p1 p2 [k3 [k4 k5 k6]]).  Pinned against cv2 4.13 (tests/golden/cv2_undistort.npz).
static inline void undistort_points_f32(const float* src_xy, int n, float fxf, float fyf, float cxf, float cyf, const float* dist, int nd, float* dst_xy)
{
    double k[14] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < nd && i < 14; i++) k[i] = (double)dist[i];
    const double fx = fxf, fy = fyf, cx = cxf, cy = cyf, ifx = 1. / fx, ify = 1. / fy;
    for (int i = 0; i < n; i++) {
        double x = src_xy[2 * i], y = src_xy[2 * i + 1];
        const double u = x, v = y;
        x = (x - cx) * ifx;
        y = (y - cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        const double xx = fx * x + 0.0 * y + cx;
        const double yy = 0.0 * x + fy * y + cy;
        const double ww = 1. / (0.0 * x + 0.0 * y + 1.0);
        dst_xy[2 * i] = (float)(xx * ww);
        dst_xy[2 * i + 1] = (float)(yy * ww);
    }
}

}  // namespace cvp


