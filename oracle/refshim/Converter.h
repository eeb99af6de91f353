// oracle/refshim/Converter.h — TEST INFRASTRUCTURE.  Stands in for the reference's include/Converter.h, which pulls in Eigen and
// g2o (absent from this image).  The only member the compiled sources use is toDescriptorVector (Frame.cpp:456, KeyFrame.cpp:69);
// its body (src/Converter.cpp:28-37: one cv::Mat row header per descriptor) is supplied by oracle/refmatch_capi.cpp.
#ifndef CONVERTER_H
#define CONVERTER_H
#include <vector>
#include <opencv2/core/core.hpp>
namespace ORB_SLAM2 {
class Converter {
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors);
};
}
#endif
