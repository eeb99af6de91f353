// orbb_cv.h — the few cv:: types the ORBextractor / ORBmatcher class surface needs.
// With OpenCV installed (the reference's normal build) the real headers are used; on machines without C++ OpenCV
// (this repo's build and GPU boxes) a minimal stand-in with the same member names and binary KeyPoint layout is
// defined, so the host mirror and its tests compile anywhere.  Define ORBB_NO_OPENCV to force the stand-in.
#ifndef ORBB_CV_H
#define ORBB_CV_H

#if !defined(ORBB_NO_OPENCV) && defined(__has_include)
#if __has_include(<opencv2/core/core.hpp>)
#include <opencv2/core/core.hpp>
#define ORBB_HAVE_OPENCV 1
#endif
#endif

#ifndef ORBB_HAVE_OPENCV
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#include <memory>
#include <vector>

#ifndef CV_8U
#define CV_8U 0
#define CV_8UC1 0
#endif

namespace cv {

struct Point2f {
    float x, y;
    Point2f() : x(0), y(0) {}
    Point2f(float x_, float y_) : x(x_), y(y_) {}
};

// same 28-byte layout as cv::KeyPoint {pt, size, angle, response, octave, class_id}
struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};

// 8-bit single-channel matrix with shared ownership (enough for images, pyramid levels and the N x 32 descriptor Mat)
class Mat {
public:
    int rows, cols;
    size_t step;
    unsigned char* data;
    Mat() : rows(0), cols(0), step(0), data(nullptr) {}
    Mat(int r, int c, int /*type*/) : rows(0), cols(0), step(0), data(nullptr) { create(r, c, CV_8U); }
    Mat(int r, int c, int /*type*/, void* ext, size_t stp = 0) : rows(r), cols(c), step(stp ? stp : (size_t)c), data((unsigned char*)ext) {}
    void create(int r, int c, int /*type*/)
    {
        if (r == rows && c == cols && own_ && step == (size_t)c) return;
        own_ = std::shared_ptr<unsigned char>(new unsigned char[(size_t)(r > 0 ? r : 0) * (c > 0 ? c : 0) + 1], std::default_delete<unsigned char[]>());
        rows = r; cols = c; step = (size_t)c; data = own_.get();
    }
    void release() { own_.reset(); rows = cols = 0; step = 0; data = nullptr; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return CV_8UC1; }
    unsigned char* ptr(int y = 0) { return data + (size_t)y * step; }
    const unsigned char* ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T> T* ptr(int y) { return (T*)(data + (size_t)y * step); }
    template <typename T> const T* ptr(int y) const { return (const T*)(data + (size_t)y * step); }
    Mat row(int y) const { Mat m(1, cols, CV_8U, (void*)ptr(y), step); m.own_ = own_; return m; }
    Mat clone() const
    {
        Mat m(rows, cols, CV_8U);
        for (int y = 0; y < rows; y++) memcpy(m.ptr(y), ptr(y), (size_t)cols);
        return m;
    }
    Mat getMat() const { return *this; }

private:
    std::shared_ptr<unsigned char> own_;
};

typedef const Mat& InputArray;
typedef Mat& OutputArray;

}  // namespace cv
#endif  // !ORBB_HAVE_OPENCV

static_assert(sizeof(cv::KeyPoint) == 28, "cv::KeyPoint must be 28 bytes: it is copied verbatim across the C ABI");

#endif  // ORBB_CV_H
