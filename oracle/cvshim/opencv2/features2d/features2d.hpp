#pragma once
#include "opencv2/core/core.hpp"
namespace cv {
static inline void FAST(const Mat& img, std::vector<KeyPoint>& kps, int threshold, bool /*nonmaxSuppression=true*/)
{
    std::vector<cvp::FastKp> v;
    cvp::fast9_16_nms(img.data, img.cols, img.rows, img.step, threshold, v);
    kps.clear();
    for (const auto& k : v) kps.push_back(KeyPoint((float)k.x, (float)k.y, 7.f, -1, (float)k.score));
}
struct KeyPointsFilter {   // only referenced by the dead ComputeKeyPointsOld (ORBextractor.cpp:865-1042)
    static void retainBest(std::vector<KeyPoint>& k, int n) {
        if (n >= 0 && (int)k.size() > n) {
            std::stable_sort(k.begin(), k.end(), [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
            k.resize(n);
        }
    }
};
}
