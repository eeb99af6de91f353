// ORBextractor.h — host-side mirror of ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:45-111) whose
// operator() runs on a B200 through the C ABI of liborbb200.so.  A maintainer drops this header + ORBextractor.cc in
// place of the reference's pair: same namespace, constructor, operator(), getters and public mvImagePyramid, so
// Frame.cc / Tracking.cc compile unchanged.
#ifndef ORBB_HOST_ORBEXTRACTOR_H
#define ORBB_HOST_ORBEXTRACTOR_H

#include <stdexcept>
#include <string>
#include <vector>
#include "orbb_cv.h"
#include "../../include/orbb200.h"

namespace ORB_SLAM2 {

// Thrown where the reference would hit a cv::Exception / assert (wrong image type) or when the device call fails;
// the library itself has no CPU fallback.
class OrbDeviceError : public std::runtime_error {
public:
    OrbDeviceError(int code_, const std::string& what_) : std::runtime_error(what_), code(code_) {}
    int code;
};

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    // include/ORBextractor.h:51-52.  The device handle is created on the first image (its size fixes the buffers)
    // or eagerly with Reserve().
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
    ~ORBextractor();
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // include/ORBextractor.h:59-61.  `mask` is ignored, as in the reference (:58).  Empty image: silent return
    // (src/ORBextractor.cpp:1056-1057); zero keypoints: descriptors released (:1076-1077).
    void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors);

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return (float)scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // include/ORBextractor.h:85 — read by Frame::ComputeStereoMatches (src/Frame.cpp:547,657-676).  Filled after every
    // operator() while SetPyramidDownload(true) (default, for drop-in fidelity); callers that moved stereo matching to
    // orbm_stereo_match (which reads the pyramid in HBM) switch it off and save the copy.
    std::vector<cv::Mat> mvImagePyramid;

    // ---- additions for the device path (not in the reference) ----
    void Reserve(int maxWidth, int maxHeight, int device = 0);   // create the handle now
    void SetPyramidDownload(bool on) { mbDownloadPyramid = on; }
    void SetDevice(int device) { mnDevice = device; }
    orbx_handle* Handle() { return mpHandle; }                    // for orbm_stereo_match(left->Handle(), right->Handle(), ...)
    int Capacity() const { return nfeatures + 64 * nlevels + 256; } // rows the octree can return at most (as sized in operator())

protected:
    void EnsureHandle(int w, int h);

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<int> umax;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;

    orbx_handle* mpHandle;
    int mnMaxW, mnMaxH, mnDevice;
    bool mbDownloadPyramid;
    std::vector<orbx_keypoint> mvStage;
};

}  // namespace ORB_SLAM2
#endif
