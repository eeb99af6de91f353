#pragma once
#include "opencv2/core/core.hpp"
namespace cv {
static inline void resize(const Mat& src, Mat& dst, Size dsize, double, double, int /*INTER_LINEAR*/)
{
    dst.create(dsize.height, dsize.width, CV_8UC1);
    cvp::resize_linear_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.cols, dst.rows, dst.step);
}
static inline void GaussianBlur(const Mat& src, Mat& dst, Size /*7x7*/, double /*2*/, double /*2*/, int /*REFLECT_101*/)
{
    Mat out; out.alloc(src.rows, src.cols);
    cvp::gaussian_blur7_u8(src.data, src.cols, src.rows, src.step, out.data, out.step);
    dst.create(src.rows, src.cols, CV_8UC1);
    for (int y = 0; y < src.rows; y++) memcpy(dst.ptr(y), out.ptr(y), src.cols);
}
}
