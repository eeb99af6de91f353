// oracle/cvshim — TEST INFRASTRUCTURE.  A minimal stand-in for the slice of the OpenCV C++ API that the reference's
// ORBextractor / ORBmatcher / Frame / KeyFrame / MapPoint / Map / KeyFrameDatabase sources and the vendored DBoW2 use, so
// that those UNMODIFIED sources compile here (there is no C++ OpenCV in this image).  The pixel arithmetic behind resize /
// GaussianBlur / FAST / fastAtan2 / undistortPoints is oracle/cv_prims.hpp, which tests pin bit-exactly to cv2 4.13.
// Small-matrix float algebra follows what OpenCV 4.x computes for CV_32F operands without FMA contraction: products summed left
// to right in float (gemm's 2..4-wide special case), norm / dot accumulated in double, M / s evaluated as M * (float)(1.0 / s).
#pragma once
#include <cstdlib>
#include <cassert>
#include <cstring>
#include <cstdint>
#include <cmath>
#include <climits>
#include <string>
#include <sstream>
#include <iostream>
#include <fstream>
#include <vector>
#include <memory>
#include <algorithm>
#include "cv_prims.hpp"

typedef unsigned char uchar;
typedef unsigned short ushort;
#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn) - 1) << 3))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32FC2 CV_MAKETYPE(CV_32F, 2)
#define CV_64FC1 CV_MAKETYPE(CV_64F, 1)
#define CV_MAT_DEPTH(t) ((t) & 7)
#define CV_MAT_CN(t) ((((t) >> 3) & 63) + 1)

static inline int cvRound(double v) { return cvp::cv_round(v); }
static inline int cvRound(float v) { return cvp::cv_round(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { return cvp::cv_floor(v); }
static inline int cvCeil(double v) { return cvp::cv_ceil(v); }

namespace cv {

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;
template <typename T> struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T _x, T _y, T _z) : x(_x), y(_y), z(_z) {}
};
typedef Point3_<float> Point3f;

struct Size { int width, height; Size() : width(0), height(0) {} Size(int w, int h) : width(w), height(h) {} };
struct Rect { int x, y, width, height; Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {} };

struct KeyPoint {
    Point2f pt; float size, angle, response; int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};

struct MatStep {
    size_t v;
    operator size_t() const { return v; }
};

static inline int depthBytes(int type) { static const int b[8] = {1, 1, 2, 2, 4, 4, 8, 2}; return b[CV_MAT_DEPTH(type)]; }

class MatExpr;

class Mat {
public:
    int flags;       // CV_MAKETYPE(depth, channels)
    int rows, cols;
    uchar* data;
    MatStep step;    // bytes between rows
    int* refcount;   // intrusive count at the head of a malloc block (pixel storage never uses operator new,
                     // see ref_capi.cpp: the bump arena is rewound after every call but pyramids persist)

    Mat() : flags(0), rows(0), cols(0), data(nullptr), refcount(nullptr) { step.v = 0; }
    Mat(Size sz, int type) : flags(0), rows(0), cols(0), data(nullptr), refcount(nullptr) { alloc(sz.height, sz.width, type); }
    Mat(int r, int c, int type) : flags(0), rows(0), cols(0), data(nullptr), refcount(nullptr) { alloc(r, c, type); }
    Mat(int r, int c, int type, void* ext, size_t stp = 0) : flags(type), rows(r), cols(c), data((uchar*)ext), refcount(nullptr)
    {
        step.v = stp ? stp : (size_t)c * elemSize();
    }
    Mat(const Mat& o) : flags(o.flags), rows(o.rows), cols(o.cols), data(o.data), step(o.step), refcount(o.refcount) { if (refcount) __sync_fetch_and_add(refcount, 1); }
    Mat& operator=(const Mat& o) {
        if (this != &o) {
            if (o.refcount) __sync_fetch_and_add(o.refcount, 1);
            drop();
            flags = o.flags; rows = o.rows; cols = o.cols; data = o.data; step = o.step; refcount = o.refcount;
        }
        return *this;
    }
    // Mat = MatExpr: cv::Mat::create keeps the buffer when size and type already match, so assigning an expression to a
    // row / column view writes through to the parent (ORBextractor.cpp:1047 relies on this for Mat::zeros).
    inline Mat& operator=(const MatExpr& e);
    ~Mat() { drop(); }
    void drop() { if (refcount && __sync_sub_and_fetch(refcount, 1) == 0) free(refcount); refcount = nullptr; }

    size_t elemSize() const { return (size_t)depthBytes(flags) * CV_MAT_CN(flags); }
    size_t elemSize1() const { return (size_t)depthBytes(flags); }
    int type() const { return flags; }
    int depth() const { return CV_MAT_DEPTH(flags); }
    int channels() const { return CV_MAT_CN(flags); }
    size_t total() const { return (size_t)rows * cols; }
    Size size() const { return Size(cols, rows); }
    bool isContinuous() const { return rows <= 1 || step.v == (size_t)cols * elemSize(); }

    void alloc(int r, int c, int type = CV_8UC1) {
        drop();
        flags = type; rows = r; cols = c; step.v = (size_t)c * elemSize();
        refcount = (int*)malloc(64 + (size_t)r * step.v + 64);
        *refcount = 1;
        data = (uchar*)refcount + 64;
    }
    void create(int r, int c, int type) { if (r != rows || c != cols || type != flags || !data) alloc(r, c, type); }
    void create(Size sz, int type) { create(sz.height, sz.width, type); }
    void release() { drop(); rows = cols = 0; data = nullptr; step.v = 0; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t step1() const { return step.v / elemSize1(); }
    template <typename T> T& at(int r, int c) { return *(T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T)); }
    template <typename T> const T& at(int r, int c) const { return *(const T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T)); }
    // at(i) on a single-row or single-column matrix
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    uchar* ptr(int r = 0) { return data + (size_t)r * step.v; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step.v; }
    template <typename T> T* ptr(int r = 0) { return (T*)(data + (size_t)r * step.v); }
    template <typename T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step.v); }
    Mat rowRange(int a, int b) const { Mat m(*this); m.data = data + (size_t)a * step.v; m.rows = b - a; return m; }
    Mat colRange(int a, int b) const { Mat m(*this); m.data = data + (size_t)a * elemSize(); m.cols = b - a; return m; }
    Mat row(int y) const { return rowRange(y, y + 1); }
    Mat col(int x) const { return colRange(x, x + 1); }
    Mat operator()(const Rect& r) const { return rowRange(r.y, r.y + r.height).colRange(r.x, r.x + r.width); }
    Mat clone() const {
        Mat m; if (empty()) return m;
        m.alloc(rows, cols, flags);
        for (int y = 0; y < rows; y++) memcpy(m.ptr(y), ptr(y), (size_t)cols * elemSize());
        return m;
    }
    void copyTo(Mat& dst) const {
        if (empty()) { dst.release(); return; }
        dst.create(rows, cols, flags);
        if (dst.data == data && dst.step.v == step.v) return;
        for (int y = 0; y < rows; y++) memmove(dst.ptr(y), ptr(y), (size_t)cols * elemSize());
    }
    void copyTo(const Mat& dstView) const { Mat d(dstView); copyTo(d); }   // copyTo(M.rowRange(..).col(..)): writes through a temporary view
    // reshape(cn): same data, new channel count (Frame.cpp:482-484 toggles an N x 2 CV_32F matrix to N x 1 CV_32FC2 and back)
    Mat reshape(int cn) const {
        Mat m(*this);
        const int total_ch = cols * channels();
        m.flags = CV_MAKETYPE(depth(), cn);
        m.cols = total_ch / cn;
        return m;
    }
    inline void convertTo(Mat& dst, int rtype) const;
    inline MatExpr t() const;
    inline double dot(const Mat& o) const;
    static inline MatExpr zeros(int r, int c, int t);
    static inline MatExpr ones(int r, int c, int t);
    static inline MatExpr eye(int r, int c, int t);

    // element access by linear index over (row, col) honouring the step: float / double matrices only
    double getd(int r, int c) const { return depth() == CV_64F ? at<double>(r, c) : (double)at<float>(r, c); }
};

// The result of an arithmetic expression.  OpenCV evaluates MatExpr lazily; here it is evaluated eagerly and only the
// assignment semantics (write-through into an existing buffer of the same shape) are kept.
class MatExpr : public Mat {
public:
    MatExpr() {}
    explicit MatExpr(const Mat& m) : Mat(m) {}
};

inline Mat& Mat::operator=(const MatExpr& e)
{
    if (data && rows == e.rows && cols == e.cols && flags == e.flags) {
        for (int y = 0; y < rows; y++) memmove(ptr(y), e.ptr(y), (size_t)cols * elemSize());
        return *this;
    }
    return *this = static_cast<const Mat&>(e);
}

inline MatExpr Mat::zeros(int r, int c, int t)
{
    MatExpr m; m.alloc(r, c, t);
    for (int y = 0; y < r; y++) memset(m.ptr(y), 0, (size_t)c * m.elemSize());
    return m;
}
inline MatExpr Mat::ones(int r, int c, int t)
{
    MatExpr m = zeros(r, c, t);
    for (int y = 0; y < r; y++)
        for (int x = 0; x < c; x++) {
            if (m.depth() == CV_32F) m.at<float>(y, x) = 1.f;
            else if (m.depth() == CV_64F) m.at<double>(y, x) = 1.0;
            else if (m.depth() == CV_8U) m.at<uchar>(y, x) = 1;
            else if (m.depth() == CV_32S) m.at<int>(y, x) = 1;
        }
    return m;
}
inline MatExpr Mat::eye(int r, int c, int t)
{
    MatExpr m = zeros(r, c, t);
    for (int i = 0; i < std::min(r, c); i++) {
        if (m.depth() == CV_32F) m.at<float>(i, i) = 1.f;
        else if (m.depth() == CV_64F) m.at<double>(i, i) = 1.0;
    }
    return m;
}
inline void Mat::convertTo(Mat& dst, int rtype) const
{
    const int dd = CV_MAT_DEPTH(rtype);
    Mat out; out.alloc(rows, cols, CV_MAKETYPE(dd, channels()));
    const int n = cols * channels();
    for (int y = 0; y < rows; y++)
        for (int x = 0; x < n; x++) {
            double v;
            switch (depth()) {
                case CV_8U: v = ptr<uchar>(y)[x]; break;
                case CV_32S: v = ptr<int>(y)[x]; break;
                case CV_32F: v = ptr<float>(y)[x]; break;
                default: v = ptr<double>(y)[x]; break;
            }
            switch (dd) {
                case CV_8U: out.ptr<uchar>(y)[x] = (uchar)std::min(255, std::max(0, cvp::cv_round(v))); break;
                case CV_32S: out.ptr<int>(y)[x] = cvp::cv_round(v); break;
                case CV_32F: out.ptr<float>(y)[x] = (float)v; break;
                default: out.ptr<double>(y)[x] = v; break;
            }
        }
    dst = out;
}
inline MatExpr Mat::t() const
{
    MatExpr m; m.alloc(cols, rows, flags);
    const size_t es = elemSize();
    for (int y = 0; y < rows; y++)
        for (int x = 0; x < cols; x++) memcpy(m.data + (size_t)x * m.step.v + (size_t)y * es, data + (size_t)y * step.v + (size_t)x * es, es);
    return m;
}
// cv::Mat::dot: both operands as flat vectors; CV_32F products accumulated in double (dotProd_32f).
inline double Mat::dot(const Mat& o) const
{
    double s = 0;
    const int n = rows * cols;
    for (int i = 0; i < n; i++) {
        const int r0 = i / cols, c0 = i % cols, r1 = i / o.cols, c1 = i % o.cols;
        if (depth() == CV_32F) s += (double)at<float>(r0, c0) * (double)o.at<float>(r1, c1);
        else s += at<double>(r0, c0) * o.at<double>(r1, c1);
    }
    return s;
}

// ---- arithmetic on CV_32F / CV_64F matrices -------------------------------------------------------------------------
template <typename F> static inline MatExpr matBinary(const Mat& a, const Mat& b, F f)
{
    assert(a.rows == b.rows && a.cols == b.cols && a.flags == b.flags);
    MatExpr m; m.alloc(a.rows, a.cols, a.flags);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) {
            if (a.depth() == CV_32F) m.at<float>(y, x) = (float)f(a.at<float>(y, x), b.at<float>(y, x));
            else m.at<double>(y, x) = f(a.at<double>(y, x), b.at<double>(y, x));
        }
    return m;
}
static inline MatExpr operator+(const Mat& a, const Mat& b) { return matBinary(a, b, [](auto p, auto q) { return p + q; }); }
static inline MatExpr operator-(const Mat& a, const Mat& b) { return matBinary(a, b, [](auto p, auto q) { return p - q; }); }
// scale by a scalar: cvtScale casts alpha to the matrix's type first (float alpha for CV_32F)
static inline MatExpr operator*(const Mat& a, double s)
{
    MatExpr m; m.alloc(a.rows, a.cols, a.flags);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) {
            if (a.depth() == CV_32F) m.at<float>(y, x) = a.at<float>(y, x) * (float)s;
            else m.at<double>(y, x) = a.at<double>(y, x) * s;
        }
    return m;
}
static inline MatExpr operator*(double s, const Mat& a) { return a * s; }
static inline MatExpr operator/(const Mat& a, double s) { return a * (1.0 / s); }
static inline MatExpr operator-(const Mat& a) { return a * -1.0; }
// matrix product; CV_32F: float products summed left to right (gemm's small-matrix path), CV_64F the same in double
static inline MatExpr operator*(const Mat& a, const Mat& b)
{
    assert(a.cols == b.rows && a.flags == b.flags);
    MatExpr m; m.alloc(a.rows, b.cols, a.flags);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < b.cols; x++) {
            if (a.depth() == CV_32F) {
                float s = a.at<float>(y, 0) * b.at<float>(0, x);
                for (int k = 1; k < a.cols; k++) s = s + a.at<float>(y, k) * b.at<float>(k, x);
                m.at<float>(y, x) = s;
            } else {
                double s = a.at<double>(y, 0) * b.at<double>(0, x);
                for (int k = 1; k < a.cols; k++) s = s + a.at<double>(y, k) * b.at<double>(k, x);
                m.at<double>(y, x) = s;
            }
        }
    return m;
}

enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4 };
// cv::norm(m): L2 with double accumulation (normL2_32f)
static inline double norm(const Mat& a)
{
    double s = 0;
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) { const double v = a.getd(y, x); s += v * v; }
    return std::sqrt(s);
}
static inline double norm(const Mat& a, const Mat& b, int normType)
{
    double s = 0;
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) {
            double d;
            if (a.depth() == CV_32F) d = (double)std::abs(a.at<float>(y, x) - b.at<float>(y, x));
            else if (a.depth() == CV_8U) d = (double)std::abs((int)a.at<uchar>(y, x) - (int)b.at<uchar>(y, x));
            else d = std::abs(a.at<double>(y, x) - b.at<double>(y, x));
            if (normType == NORM_L1) s += d; else if (normType == NORM_INF) s = std::max(s, d); else s += d * d;
        }
    return normType == NORM_L2 ? std::sqrt(s) : s;
}

// cv::Mat_<T>(r, c) << a, b, c  (Frame.cpp:794, KeyFrame.cpp:91,761)
template <typename T> struct MatTypeOf;
template <> struct MatTypeOf<float> { enum { value = CV_32F }; };
template <> struct MatTypeOf<double> { enum { value = CV_64F }; };
template <> struct MatTypeOf<uchar> { enum { value = CV_8U }; };
template <> struct MatTypeOf<int> { enum { value = CV_32S }; };
template <typename T> class Mat_ : public Mat {
public:
    Mat_() {}
    Mat_(int r, int c) : Mat(r, c, MatTypeOf<T>::value) {}
    T& operator()(int r, int c) { return this->template at<T>(r, c); }
};
template <typename T> class MatCommaInitializer_ {
public:
    Mat_<T> m; int i;
    MatCommaInitializer_(const Mat_<T>& _m) : m(_m), i(0) {}
    template <typename V> MatCommaInitializer_& operator,(V v) { m.template at<T>(i / m.cols, i % m.cols) = (T)v; i++; return *this; }
    operator Mat() const { return m; }
    operator Mat_<T>() const { return m; }
};
template <typename T, typename V> static inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, V v)
{
    MatCommaInitializer_<T> ci(m);
    return (ci, v);
}

class _InputArray {
public:
    const Mat* m;
    _InputArray(const Mat& _m) : m(&_m) {}
    bool empty() const { return m->empty(); }
    Mat getMat() const { return *m; }
};
class _OutputArray {
public:
    Mat* m;
    _OutputArray(Mat& _m) : m(&_m) {}
    void release() const { m->release(); }
    void create(int r, int c, int t) const { m->create(r, c, t); }
    Mat getMat() const { return *m; }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

static inline float fastAtan2(float y, float x) { return cvp::fast_atan2(y, x); }

enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4,
       BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
enum { INTER_NEAREST = 0, INTER_LINEAR = 1 };

static inline void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int /*borderType*/)
{
    const int w = src.cols, h = src.rows, W = w + left + right, H = h + top + bottom;
    std::vector<uchar> tmp((size_t)W * H);
    for (int y = 0; y < H; y++) {
        const uchar* S = src.ptr(cvp::reflect101(y - top, h));
        for (int x = 0; x < W; x++) tmp[(size_t)y * W + x] = S[cvp::reflect101(x - left, w)];
    }
    dst.create(H, W, CV_8UC1);
    for (int y = 0; y < H; y++) memcpy(dst.ptr(y), &tmp[(size_t)y * W], W);
}

// cv::undistortPoints(src, dst, K, dist, R = empty, P = K) on an N x 1 CV_32FC2 matrix (Frame.cpp:483, :515); src may be dst.
static inline void undistortPoints(const Mat& src, Mat& dst, const Mat& K, const Mat& dist, const Mat& /*R*/, const Mat& /*P == K*/)
{
    const int n = src.rows * src.cols;
    std::vector<float> in(2 * (size_t)n + 2), out(2 * (size_t)n + 2), dc;
    for (int i = 0; i < n; i++) { const float* p = src.ptr<float>(src.cols == 1 ? i : 0) + (src.cols == 1 ? 0 : 2 * i); in[2 * i] = p[0]; in[2 * i + 1] = p[1]; }
    for (int i = 0; i < dist.rows * dist.cols; i++) dc.push_back(dist.rows == 1 ? dist.at<float>(0, i) : dist.at<float>(i, 0));
    cvp::undistort_points_f32(in.data(), n, K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2), dc.data(), (int)dc.size(), out.data());
    Mat d; d.alloc(src.rows, src.cols, CV_32FC2);
    for (int i = 0; i < n; i++) { float* p = d.ptr<float>(d.cols == 1 ? i : 0) + (d.cols == 1 ? 0 : 2 * i); p[0] = out[2 * i]; p[1] = out[2 * i + 1]; }
    dst = d;
}

// cv::FileStorage / FileNode: only named by TemplatedVocabulary::save / load (YAML), which nothing here calls
// (the reference loads ORBvoc.txt through loadFromTextFile).  Present so the class template instantiates.
class FileNode {
public:
    FileNode operator[](const char*) const { return FileNode(); }
    FileNode operator[](const std::string&) const { return FileNode(); }
    FileNode operator[](int) const { return FileNode(); }
    size_t size() const { return 0; }
    operator int() const { return 0; }
    operator float() const { return 0.f; }
    operator double() const { return 0.0; }
    operator std::string() const { return std::string(); }
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage() {}
    FileStorage(const char*, int) {}
    FileStorage(const std::string&, int) {}
    bool isOpened() const { return false; }
    void release() {}
    FileNode operator[](const char*) const { return FileNode(); }
    FileNode operator[](const std::string&) const { return FileNode(); }
};
template <typename T> static inline FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }

}  // namespace cv
