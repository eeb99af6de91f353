// ORBextractor.cc — host mirror of ORB_SLAM2::ORBextractor over liborbb200's C ABI (see ORBextractor.h).
#include "ORBextractor.h"
#include <math.h>
#include <string.h>

namespace ORB_SLAM2 {

static void check(int rc, const char* where)
{
    if (rc == ORBB_OK) return;
    std::string msg = std::string(where) + ": " + orbb200_strerror(rc);
    if (rc == ORBB_E_CUDA) msg += std::string(" — ") + orbb200_last_cuda_error();
    throw OrbDeviceError(rc, msg);
}

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST),
      mpHandle(nullptr), mnMaxW(0), mnMaxH(0), mnDevice(0), mbDownloadPyramid(true)
{
    // The scale tables are needed by Tracking/Frame before the first image arrives (src/Tracking.cc:141-149,
    // src/Frame.cc:70-77), so they are evaluated here with the reference's own float/double sequence
    // (src/ORBextractor.cpp:417-433); orbx_create evaluates the same sequence for the device and a mismatch is an error.
    mvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvScaleFactor[0] = 1.0f;
    mvLevelSigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        mvScaleFactor[i] = (float)(mvScaleFactor[i - 1] * scaleFactor);
        mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
    }
    mvInvScaleFactor.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels);
    for (int i = 0; i < nlevels; i++) {
        mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
        mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
    }
    mvImagePyramid.resize(nlevels);
    mnFeaturesPerLevel.assign(nlevels, 0);
    umax.assign(16, 0);
}

ORBextractor::~ORBextractor()
{
    if (mpHandle) orbx_destroy(mpHandle);
}

void ORBextractor::Reserve(int maxWidth, int maxHeight, int device)
{
    mnDevice = device;
    EnsureHandle(maxWidth, maxHeight);
}

void ORBextractor::EnsureHandle(int w, int h)
{
    if (mpHandle && w <= mnMaxW && h <= mnMaxH) return;
    if (mpHandle) { orbx_destroy(mpHandle); mpHandle = nullptr; }
    orbx_config cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.nfeatures = nfeatures; cfg.scale_factor = (float)scaleFactor; cfg.nlevels = nlevels;
    cfg.ini_th_fast = iniThFAST; cfg.min_th_fast = minThFAST;
    cfg.max_width = w > mnMaxW ? w : mnMaxW; cfg.max_height = h > mnMaxH ? h : mnMaxH;
    cfg.max_batch = 2; cfg.device = mnDevice; cfg.blur_legacy = 0;      // two frames: a stereo pair as one batch (orbm_stereo_frame_create)
    check(orbx_create(&cfg, &mpHandle), "orbx_create");
    mnMaxW = cfg.max_width; mnMaxH = cfg.max_height;
    std::vector<float> sc(nlevels), isc(nlevels), s2(nlevels), is2(nlevels);
    check(orbx_get_tables(mpHandle, sc.data(), isc.data(), s2.data(), is2.data(), mnFeaturesPerLevel.data(), umax.data()), "orbx_get_tables");
    if (memcmp(sc.data(), mvScaleFactor.data(), sizeof(float) * nlevels) || memcmp(isc.data(), mvInvScaleFactor.data(), sizeof(float) * nlevels) ||
        memcmp(s2.data(), mvLevelSigma2.data(), sizeof(float) * nlevels) || memcmp(is2.data(), mvInvLevelSigma2.data(), sizeof(float) * nlevels))
        throw OrbDeviceError(ORBB_E_STATE, "host and device scale tables differ");
}

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints, cv::OutputArray _descriptors)
{
    if (_image.empty()) return;                                            // src/ORBextractor.cpp:1056-1057
    cv::Mat image = _image.getMat();
    if (image.type() != CV_8UC1) throw OrbDeviceError(ORBB_E_INVALID, "image.type() == CV_8UC1 (src/ORBextractor.cpp:1060)");
    EnsureHandle(image.cols, image.rows);

    int capacity = nfeatures + 64 * nlevels + 256;                         // the octree may return slightly more than nfeatures
    int n = 0, rc;
    std::vector<unsigned char> desc;
    for (;;) {
        mvStage.resize(capacity);
        desc.resize((size_t)capacity * ORBB_DESC_BYTES);
        rc = orbx_extract(mpHandle, image.ptr(0), image.cols, image.rows, (size_t)image.step, mvStage.data(), desc.data(), capacity, &n);
        if (rc == ORBB_E_CAPACITY && n > capacity) { capacity = n + 16; continue; }
        break;
    }
    check(rc, "orbx_extract");

    if (n == 0) _descriptors.release();                                    // :1076-1077
    else {
        _descriptors.create(n, 32, CV_8U);                                 // :1080
        cv::Mat d = _descriptors.getMat();
        for (int i = 0; i < n; i++) memcpy(d.ptr(i), &desc[(size_t)i * ORBB_DESC_BYTES], ORBB_DESC_BYTES);
    }
    _keypoints.resize(n);                                                   // :1084-1085 (clear + reserve + insert)
    if (n) memcpy((void*)_keypoints.data(), mvStage.data(), sizeof(orbx_keypoint) * (size_t)n);

    if (mbDownloadPyramid) {                                                // mvImagePyramid, include/ORBextractor.h:85
        for (int l = 0; l < nlevels; l++) {
            int w = 0, h = 0;
            check(orbx_pyramid_level(mpHandle, 0, l, nullptr, 0, &w, &h), "orbx_pyramid_level");
            mvImagePyramid[l].create(h, w, CV_8U);
            check(orbx_pyramid_level(mpHandle, 0, l, mvImagePyramid[l].ptr(0), (size_t)mvImagePyramid[l].step, &w, &h), "orbx_pyramid_level");
        }
    }
}

}  // namespace ORB_SLAM2
