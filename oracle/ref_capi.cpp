// oracle/ref_capi.cpp — TEST INFRASTRUCTURE.  C API over the reference's own ORB_SLAM2::ORBextractor
// (compiled unmodified from /root/reference/ORB-SLAM2/src/ORBextractor.cpp, see oracle/Makefile).
//
// The reference sorts octree nodes by (count, heap address) (ORBextractor.cpp:690).  This library
// replaces operator new with a monotonic bump allocator (linked -Bsymbolic so only this .so sees it):
// addresses then increase with creation order, which is exactly the canonical tie-break the oracle
// and the CUDA path implement.
#include <new>
#include <cstdio>
#include "ORBextractor.h"

namespace {
struct Arena {
    static const size_t CHUNK = size_t(1) << 30;
    char* base = nullptr; size_t used = 0;
    void* take(size_t n) {
        if (!base) base = (char*)malloc(CHUNK);
        n = (n + 15) & ~size_t(15);
        if (used + n > CHUNK) { fprintf(stderr, "orb_ref arena exhausted\n"); abort(); }
        void* p = base + used; used += n; return p;
    }
};
thread_local Arena g_arena;   // one arena per calling thread: the timed CPU baseline runs one extractor per thread
}

void* operator new(size_t n) { return g_arena.take(n); }
void* operator new[](size_t n) { return g_arena.take(n); }
void operator delete(void*) noexcept {}
void operator delete[](void*) noexcept {}
void operator delete(void*, size_t) noexcept {}
void operator delete[](void*, size_t) noexcept {}

extern "C" {
void* ref_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh)
{
    return new ORB_SLAM2::ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
}
void ref_destroy(void*) {}
int ref_extract(void* h, const uint8_t* img, int w, int h_, long stride, void* kps_out, uint8_t* desc_out, int cap)
{
    ORB_SLAM2::ORBextractor* ex = (ORB_SLAM2::ORBextractor*)h;
    const size_t mark = g_arena.used;          // everything allocated below dies with this call
    int n;
    {
        cv::Mat image(h_, w, CV_8UC1, (void*)img, (size_t)stride), desc;
        std::vector<cv::KeyPoint> kps;
        (*ex)(image, cv::Mat(), kps, desc);
        n = (int)kps.size();
        static_assert(sizeof(cv::KeyPoint) == 28, "cv::KeyPoint layout");
        int m = n < cap ? n : cap;
        if (m) { memcpy(kps_out, kps.data(), (size_t)m * 28); for (int i = 0; i < m; i++) memcpy(desc_out + (size_t)i * 32, desc.ptr(i), 32); }
    }
    g_arena.used = mark;
    return n;
}
void ref_level_size(void* h, int level, int* w, int* hh)
{
    ORB_SLAM2::ORBextractor* ex = (ORB_SLAM2::ORBextractor*)h;
    *w = ex->mvImagePyramid[level].cols; *hh = ex->mvImagePyramid[level].rows;
}
void ref_level_image(void* h, int level, uint8_t* dst)
{
    ORB_SLAM2::ORBextractor* ex = (ORB_SLAM2::ORBextractor*)h;
    const cv::Mat& m = ex->mvImagePyramid[level];
    for (int y = 0; y < m.rows; y++) memcpy(dst + (size_t)y * m.cols, m.ptr(y), m.cols);
}
}
