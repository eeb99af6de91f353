// oracle/cvshim — TEST INFRASTRUCTURE.  A minimal stand-in for the slice of the OpenCV C++ API that
// the reference's ORBextractor.cpp / ORBextractor.h use, so that the UNMODIFIED reference sources
// compile here (there is no C++ OpenCV in this image).  The pixel arithmetic behind resize /
// GaussianBlur / FAST / fastAtan2 is oracle/cv_prims.hpp, which tests pin bit-exactly to cv2 4.13.
#pragma once
#include <cstdlib>
#include <cassert>
#include <cstring>
#include <cstdint>
#include <cmath>
#include <vector>
#include <memory>
#include <algorithm>
#include "cv_prims.hpp"

typedef unsigned char uchar;
#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5

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

struct MatZeros { int rows, cols, type; };

class Mat {
public:
    int rows, cols;
    uchar* data;
    MatStep step;
    int* refcount;   // intrusive count at the head of a malloc block (pixel storage never uses operator new,
                     // see ref_capi.cpp: the bump arena is rewound after every call but pyramids persist)

    Mat() : rows(0), cols(0), data(nullptr), refcount(nullptr) { step.v = 0; }
    Mat(Size sz, int /*type*/) : refcount(nullptr) { alloc(sz.height, sz.width); }
    Mat(int r, int c, int /*type*/) : refcount(nullptr) { alloc(r, c); }
    Mat(int r, int c, int /*type*/, void* ext, size_t stp) : rows(r), cols(c), data((uchar*)ext), refcount(nullptr) { step.v = stp; }
    Mat(const Mat& o) : rows(o.rows), cols(o.cols), data(o.data), step(o.step), refcount(o.refcount) { if (refcount) ++*refcount; }
    Mat& operator=(const Mat& o) {
        if (this != &o) {
            if (o.refcount) ++*o.refcount;
            drop();
            rows = o.rows; cols = o.cols; data = o.data; step = o.step; refcount = o.refcount;
        }
        return *this;
    }
    ~Mat() { drop(); }
    void drop() { if (refcount && --*refcount == 0) free(refcount); refcount = nullptr; }

    void alloc(int r, int c) {
        drop();
        rows = r; cols = c; step.v = (size_t)c;
        refcount = (int*)malloc(64 + (size_t)r * c + 64);
        *refcount = 1;
        data = (uchar*)refcount + 64;
    }
    void create(int r, int c, int /*type*/) { if (r != rows || c != cols || !data) alloc(r, c); }
    void release() { drop(); rows = cols = 0; data = nullptr; step.v = 0; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return CV_8UC1; }
    size_t step1() const { return step.v; }
    template <typename T> T& at(int r, int c) { return *(T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T)); }
    template <typename T> const T& at(int r, int c) const { return *(const T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T)); }
    uchar* ptr(int r = 0) { return data + (size_t)r * step.v; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step.v; }
    Mat rowRange(int a, int b) const { Mat m(*this); m.data = data + (size_t)a * step.v; m.rows = b - a; return m; }
    Mat colRange(int a, int b) const { Mat m(*this); m.data = data + a; m.cols = b - a; return m; }
    Mat operator()(const Rect& r) const { return rowRange(r.y, r.y + r.height).colRange(r.x, r.x + r.width); }
    Mat clone() const {
        Mat m; if (empty()) return m;
        m.alloc(rows, cols);
        for (int y = 0; y < rows; y++) memcpy(m.ptr(y), ptr(y), cols);
        return m;
    }
    static MatZeros zeros(int r, int c, int t) { return MatZeros{r, c, t}; }
    // Mat = MatExpr(zeros): cv::Mat::create keeps the buffer when size and type already match, so a
    // row-range view is zero-filled in place (ORBextractor.cpp:1047 relies on this).
    Mat& operator=(const MatZeros& z) {
        create(z.rows, z.cols, z.type);
        for (int y = 0; y < rows; y++) memset(ptr(y), 0, cols);
        return *this;
    }
};

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

}  // namespace cv
