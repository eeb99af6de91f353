// ORBmatcher.h — host-side mirror of the Hamming core of ORB_SLAM2::ORBmatcher (reference include/ORBmatcher.h:37-141)
// over liborbb200's C ABI.  The reference's Search* members take Frame / KeyFrame / MapPoint objects, which stay in the
// reference; what moves to the device is their inner loop: GetFeaturesInArea + DescriptorDistance + top-1/top-2
// bookkeeping.  This header offers the same constants and DescriptorDistance, plus RAII wrappers for the device-side
// Frame image and descriptor database, so the bodies of ORBmatcher.cc shrink to "fill queries -> one call -> copy the
// result into mvpMapPoints" (INTEGRATION.md shows the patch).
#ifndef ORBB_HOST_ORBMATCHER_H
#define ORBB_HOST_ORBMATCHER_H

#include <stdint.h>
#include <stdexcept>
#include <string>
#include <vector>
#include "orbb_cv.h"
#include "ORBextractor.h"
#include "../../include/orbb200.h"

namespace ORB_SLAM2 {

class MatcherDeviceError : public std::runtime_error {
public:
    MatcherDeviceError(int code_, const std::string& what_) : std::runtime_error(what_), code(code_) {}
    int code;
};

inline void orbm_check(int rc, const char* where)
{
    if (rc == ORBB_OK) return;
    std::string msg = std::string(where) + ": " + orbb200_strerror(rc);
    if (rc == ORBB_E_CUDA) msg += std::string(" — ") + orbb200_last_cuda_error();
    throw MatcherDeviceError(rc, msg);
}

// Frame::UndistortKeyPoints / ComputeImageBounds (src/Frame.cpp:462-536).  K4 = {fx, fy, cx, cy}; distCoef = mDistCoef (4, 5, 8 or 12 floats).
inline void UndistortKeyPoints(const std::vector<cv::KeyPoint>& mvKeys, const float K4[4], const std::vector<float>& distCoef,
                               std::vector<cv::KeyPoint>& mvKeysUn, int device = 0)
{
    mvKeysUn.resize(mvKeys.size());
    orbm_check(orbm_undistort_keypoints((const orbx_keypoint*)mvKeys.data(), (int)mvKeys.size(), K4, distCoef.data(), (int)distCoef.size(),
                                        (orbx_keypoint*)mvKeysUn.data(), device), "orbm_undistort_keypoints");
}
inline void ComputeImageBounds(int cols, int rows, const float K4[4], const std::vector<float>& distCoef, float& mnMinX, float& mnMinY,
                               float& mnMaxX, float& mnMaxY, int device = 0)
{
    float b[4];
    orbm_check(orbm_image_bounds(cols, rows, K4, distCoef.data(), (int)distCoef.size(), b, device), "orbm_image_bounds");
    mnMinX = b[0]; mnMinY = b[1]; mnMaxX = b[2]; mnMaxY = b[3];
}

// Device image of a Frame's matching state: mvKeysUn, mDescriptors and the 64x48 grid (src/Frame.cc:241-257).
class DeviceFrame {
public:
    DeviceFrame(const std::vector<cv::KeyPoint>& keysUn, const cv::Mat& descriptors, float minX, float minY, float maxX, float maxY, int device = 0)
        : f_(nullptr), n_((int)keysUn.size())
    {
        std::vector<uint8_t> d((size_t)n_ * 32);
        for (int i = 0; i < n_; i++) memcpy(&d[(size_t)i * 32], descriptors.ptr(i), 32);
        orbm_check(orbm_frame_create(&f_, (const orbx_keypoint*)keysUn.data(), d.data(), n_, minX, minY, maxX, maxY, device), "orbm_frame_create");
    }
    // Frame::Frame, monocular constructor (src/Frame.cpp:171-224) as one device call (orbm_mono_frame_create): ExtractORB(0, imGray),
    // UndistortKeyPoints and AssignFeaturesToGrid.  distCoef empty or distCoef[0] == 0: mvKeysUn = mvKeys (:465-469).
    DeviceFrame(ORBextractor& extractor, const cv::Mat& imGray, const float K4[4], const std::vector<float>& distCoef, float minX, float minY, float maxX,
                float maxY, std::vector<cv::KeyPoint>& mvKeys, std::vector<cv::KeyPoint>& mvKeysUn, cv::Mat& mDescriptors)
        : f_(nullptr), n_(0)
    {
        extractor.Reserve(imGray.cols, imGray.rows);
        const int cap = extractor.Capacity();
        int n = 0;
        mvKeys.resize(cap); mvKeysUn.resize(cap); mDescriptors.create(cap, 32, CV_8U);
        orbm_check(orbm_mono_frame_create(extractor.Handle(), imGray.ptr(0), imGray.cols, imGray.rows, imGray.step, distCoef.empty() ? nullptr : K4,
                                          distCoef.empty() ? nullptr : distCoef.data(), (int)distCoef.size(), minX, minY, maxX, maxY, &f_,
                                          (orbx_keypoint*)mvKeys.data(), (orbx_keypoint*)mvKeysUn.data(), mDescriptors.ptr(0), &n, cap), "orbm_mono_frame_create");
        n_ = n;
        mvKeys.resize(n); mvKeysUn.resize(n); mDescriptors.rows = n;
    }
    // Frame::Frame, stereo constructor (src/Frame.cpp:61-115) as one device call (orbm_stereo_frame_create): both images are extracted
    // as one batch of `extractor`'s handle (it plays mpORBextractorLeft and ...Right), ComputeStereoMatches and AssignFeaturesToGrid run
    // on the results in device memory.  Fills every vector the constructor fills; mvKeysUn = mvKeys (rectified input, :467-472).
    // SearchByProjection / SearchByProjectionLastFrame with mvuRight == StoredURight() reuse the device copy of mvuRight.
    static const std::vector<float>* StoredURight() { static const std::vector<float> tag; return &tag; }
    DeviceFrame(ORBextractor& extractor, const cv::Mat& imLeft, const cv::Mat& imRight, float mbf, float mb, float minX, float minY, float maxX,
                float maxY, std::vector<cv::KeyPoint>& mvKeys, cv::Mat& mDescriptors, std::vector<cv::KeyPoint>& mvKeysRight, cv::Mat& mDescriptorsRight,
                std::vector<float>& mvuRight, std::vector<float>& mvDepth)
        : f_(nullptr), n_(0)
    {
        extractor.Reserve(imLeft.cols, imLeft.rows);
        const int cap = extractor.Capacity();
        int nL = 0, nR = 0;
        mvKeys.resize(cap); mvKeysRight.resize(cap); mvuRight.assign(cap, -1.0f); mvDepth.assign(cap, -1.0f);
        mDescriptors.create(cap, 32, CV_8U); mDescriptorsRight.create(cap, 32, CV_8U);
        orbm_check(orbm_stereo_frame_create(extractor.Handle(), imLeft.ptr(0), imRight.ptr(0), imLeft.cols, imLeft.rows, imLeft.step, mbf, mb, minX, minY,
                                            maxX, maxY, &f_, (orbx_keypoint*)mvKeys.data(), mDescriptors.ptr(0), &nL, (orbx_keypoint*)mvKeysRight.data(),
                                            mDescriptorsRight.ptr(0), &nR, cap, mvuRight.data(), mvDepth.data()), "orbm_stereo_frame_create");
        n_ = nL;
        mvKeys.resize(nL); mvKeysRight.resize(nR); mvuRight.resize(nL); mvDepth.resize(nL);
        mDescriptors.rows = nL; mDescriptorsRight.rows = nR;
    }
    ~DeviceFrame() { if (f_) orbm_frame_destroy(f_); }
    DeviceFrame(const DeviceFrame&) = delete;
    DeviceFrame& operator=(const DeviceFrame&) = delete;
    const orbm_frame* get() const { return f_; }
    int size() const { return n_; }

private:
    orbm_frame* f_;
    int n_;
};

// One shard of a keyframe-descriptor database for relocalisation / loop-candidate brute-force matching.
class DeviceDescriptorDB {
public:
    DeviceDescriptorDB(int64_t capacityRows, int64_t rowOffset = 0, int device = 0) : db_(nullptr)
    {
        orbm_check(orbm_db_create(&db_, capacityRows, rowOffset, device), "orbm_db_create");
    }
    ~DeviceDescriptorDB() { if (db_) orbm_db_destroy(db_); }
    DeviceDescriptorDB(const DeviceDescriptorDB&) = delete;
    DeviceDescriptorDB& operator=(const DeviceDescriptorDB&) = delete;
    void Append(const cv::Mat& descriptors)
    {
        std::vector<uint8_t> d((size_t)descriptors.rows * 32);
        for (int i = 0; i < descriptors.rows; i++) memcpy(&d[(size_t)i * 32], descriptors.ptr(i), 32);
        orbm_check(orbm_db_append(db_, d.data(), descriptors.rows), "orbm_db_append");
    }
    int64_t Rows() const { return orbm_db_rows(db_); }
    // best distance, GLOBAL row of the best (first on ties), second-best distance per query row
    void Top2(const cv::Mat& queries, std::vector<int32_t>& d1, std::vector<int64_t>& idx1, std::vector<int32_t>& d2) const
    {
        const int nq = queries.rows;
        std::vector<uint8_t> q((size_t)nq * 32);
        for (int i = 0; i < nq; i++) memcpy(&q[(size_t)i * 32], queries.ptr(i), 32);
        d1.resize(nq); idx1.resize(nq); d2.resize(nq);
        orbm_check(orbm_bf_top2(db_, q.data(), nq, d1.data(), idx1.data(), d2.data()), "orbm_bf_top2");
    }
    orbm_db* get() const { return db_; }

private:
    orbm_db* db_;
};

// Frame::ComputeStereoMatches (src/Frame.cpp:540-748) on the device: reads both extractors' pyramids in HBM (their last operator()
// call), so SetPyramidDownload(false) is enough for stereo frames.  mb = the member value at call time (0 in the reference's stereo
// constructor).  Fills mvuRight / mvDepth (-1 = no match).
inline void ComputeStereoMatches(ORBextractor& left, ORBextractor& right, const std::vector<cv::KeyPoint>& mvKeys, const cv::Mat& mDescriptors,
                                 const std::vector<cv::KeyPoint>& mvKeysRight, const cv::Mat& mDescriptorsRight, float mbf, float mb,
                                 std::vector<float>& mvuRight, std::vector<float>& mvDepth)
{
    const int nL = (int)mvKeys.size(), nR = (int)mvKeysRight.size();
    std::vector<uint8_t> dL((size_t)nL * 32), dR((size_t)nR * 32);
    for (int i = 0; i < nL; i++) memcpy(&dL[(size_t)i * 32], mDescriptors.ptr(i), 32);
    for (int i = 0; i < nR; i++) memcpy(&dR[(size_t)i * 32], mDescriptorsRight.ptr(i), 32);
    mvuRight.assign(nL, -1.0f); mvDepth.assign(nL, -1.0f);
    orbm_check(orbm_stereo_match(left.Handle(), right.Handle(), (const orbx_keypoint*)mvKeys.data(), dL.data(), nL,
                                 (const orbx_keypoint*)mvKeysRight.data(), dR.data(), nR, mbf, mb, mvuRight.data(), mvDepth.data()), "orbm_stereo_match");
}

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

    // include/ORBmatcher.h:45, src/ORBmatcher.cpp:1838-1854.  Routed to the device like everything else in this mirror (no host
    // arithmetic), so ONE pair costs a kernel launch and three copies: a caller that keeps the reference's per-pair loops
    // (Frame::ComputeStereoMatches, src/Frame.cpp:632; MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cpp:345) would run
    // ~10^4 times slower than the reference.  Those two loops are whole entry points here: ComputeStereoMatches() below
    // (orbm_stereo_match) and orbm_distinctive_descriptors; any other set of pairs goes through DescriptorDistances() in one call.
    [[deprecated("one device round trip per pair: batch with ORBmatcher::DescriptorDistances, or use ComputeStereoMatches / orbm_distinctive_descriptors")]]
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b, int device = 0)
    {
        int32_t d = -1;
        orbm_check(orbm_distance_pairs(a.ptr(0), b.ptr(0), 1, &d, device), "orbm_distance_pairs");
        return d;
    }
    // n pairs (a[i], b[i]) of 32-byte descriptors in one launch
    static void DescriptorDistances(const uint8_t* a, const uint8_t* b, int n, int32_t* out, int device = 0)
    {
        orbm_check(orbm_distance_pairs(a, b, n, out, device), "orbm_distance_pairs");
    }

    static const float* URightArg(const std::vector<float>* v) { return !v ? nullptr : (v == DeviceFrame::StoredURight() ? ORBM_URIGHT_OF_FRAME : v->data()); }
    // src/ORBmatcher.cpp:63-155 after the caller filled one orbm_proj_query per visible map point.
    int SearchByProjection(const DeviceFrame& F, const std::vector<float>* mvuRight, const std::vector<orbm_proj_query>& q,
                           std::vector<int32_t>& frameMapPointIds, std::vector<int32_t>& frameMapPointHasObs) const
    {
        int n = 0;
        orbm_check(orbm_search_by_projection_points(F.get(), URightArg(mvuRight), q.data(), (int)q.size(), mfNNratio,
                                                    frameMapPointIds.data(), frameMapPointHasObs.data(), &n), "orbm_search_by_projection_points");
        return n;
    }
    // src/ORBmatcher.cpp:1506-1659 (Current vs Last frame) after the projection step.
    int SearchByProjectionLastFrame(const DeviceFrame& Current, const std::vector<float>* mvuRight, const std::vector<orbm_proj_query>& q,
                                    std::vector<int32_t>& frameMapPointIds, std::vector<int32_t>& frameMapPointHasObs) const
    {
        int n = 0;
        orbm_check(orbm_search_by_projection_frame(Current.get(), URightArg(mvuRight), q.data(), (int)q.size(),
                                                   mbCheckOrientation ? 1 : 0, frameMapPointIds.data(), frameMapPointHasObs.data(), &n),
                   "orbm_search_by_projection_frame");
        return n;
    }
    // src/ORBmatcher.cpp:474-590
    int SearchForInitialization(const DeviceFrame& F1, const DeviceFrame& F2, std::vector<cv::Point2f>& vbPrevMatched,
                                std::vector<int>& vnMatches12, int windowSize = 10) const
    {
        int n = 0;
        vnMatches12.assign(F1.size(), -1);
        static_assert(sizeof(cv::Point2f) == 8, "Point2f layout");
        orbm_check(orbm_search_for_initialization(F1.get(), F2.get(), (float*)vbPrevMatched.data(), vnMatches12.data(), windowSize, mfNNratio,
                                                  mbCheckOrientation ? 1 : 0, &n), "orbm_search_for_initialization");
        return n;
    }

    // src/ORBmatcher.cpp:1661-1789 (relocalisation): frameMapPointIds[i] >= 0 <=> CurrentFrame.mvpMapPoints[i] is set.
    int SearchByProjectionReloc(const DeviceFrame& Current, const std::vector<orbm_proj_query>& q, int ORBdist,
                                std::vector<int32_t>& frameMapPointIds) const
    {
        int n = 0;
        orbm_check(orbm_search_by_projection_reloc(Current.get(), q.data(), (int)q.size(), ORBdist, mbCheckOrientation ? 1 : 0,
                                                   frameMapPointIds.data(), &n), "orbm_search_by_projection_reloc");
        return n;
    }
    // src/ORBmatcher.cpp:352-472 (loop closing, Sim3-projected points): vpMatchedIds[i] >= 0 <=> vpMatched[i] is set.
    int SearchByProjectionLoop(const DeviceFrame& KF, const std::vector<orbm_proj_query>& q, std::vector<int32_t>& vpMatchedIds) const
    {
        int n = 0;
        orbm_check(orbm_search_by_projection_loop(KF.get(), q.data(), (int)q.size(), vpMatchedIds.data(), &n), "orbm_search_by_projection_loop");
        return n;
    }
    // The per-point search of Fuse (src/ORBmatcher.cpp:962-1115 with the chi-square gate, :1119-1251 without) and of each direction
    // of SearchBySim3 (:1253-1500, maxDist = TH_HIGH): bestIdx[i] = feature of KF selected for point i, or -1.
    void FuseBest(const DeviceFrame& KF, const std::vector<orbm_proj_query>& q, const std::vector<float>* mvuRight,
                  const std::vector<float>* mvInvLevelSigma2, int maxDist, std::vector<int32_t>& bestIdx) const
    {
        bestIdx.assign(q.size(), -1);
        orbm_check(orbm_fuse_best(KF.get(), q.data(), (int)q.size(), mvuRight ? mvuRight->data() : nullptr,
                                  mvInvLevelSigma2 ? mvInvLevelSigma2->data() : nullptr, maxDist, bestIdx.data(), nullptr), "orbm_fuse_best");
    }
    // src/ORBmatcher.cpp:205-348 and :604-743 on flattened FeatureVectors (node ids ascending, CSR offsets, feature indices) as
    // orbv_compute_bow returns them.  hasMapPoint[i] != 0 <=> the feature's MapPoint exists and is not bad.
    struct FeatVec { std::vector<int32_t> nodes, offsets, features; };
    int SearchByBoW(const cv::Mat& descKF, const std::vector<float>& angleKF, const std::vector<uint8_t>& hasMapPointKF, const FeatVec& fvKF,
                    const cv::Mat& descF, const std::vector<float>& angleF, const FeatVec& fvF, std::vector<int32_t>& matchFtoKF, int device = 0) const
    {
        int n = 0;
        matchFtoKF.assign(descF.rows, -1);
        orbm_check(orbm_search_by_bow(descKF.ptr(0), angleKF.data(), hasMapPointKF.data(), descKF.rows, fvKF.nodes.data(), fvKF.offsets.data(),
                                      fvKF.features.data(), (int)fvKF.nodes.size(), descF.ptr(0), angleF.data(), descF.rows, fvF.nodes.data(),
                                      fvF.offsets.data(), fvF.features.data(), (int)fvF.nodes.size(), mfNNratio, mbCheckOrientation ? 1 : 0,
                                      matchFtoKF.data(), &n, device), "orbm_search_by_bow");
        return n;
    }
    int SearchByBoW(const cv::Mat& desc1, const std::vector<float>& angle1, const std::vector<uint8_t>& hasMapPoint1, const FeatVec& fv1,
                    const cv::Mat& desc2, const std::vector<float>& angle2, const std::vector<uint8_t>& hasMapPoint2, const FeatVec& fv2,
                    std::vector<int32_t>& match12, int device = 0) const
    {
        int n = 0;
        match12.assign(desc1.rows, -1);
        orbm_check(orbm_search_by_bow_kf(desc1.ptr(0), angle1.data(), hasMapPoint1.data(), desc1.rows, fv1.nodes.data(), fv1.offsets.data(),
                                         fv1.features.data(), (int)fv1.nodes.size(), desc2.ptr(0), angle2.data(), hasMapPoint2.data(), desc2.rows,
                                         fv2.nodes.data(), fv2.offsets.data(), fv2.features.data(), (int)fv2.nodes.size(), mfNNratio,
                                         mbCheckOrientation ? 1 : 0, match12.data(), &n, device), "orbm_search_by_bow_kf");
        return n;
    }

    // src/ORBmatcher.cpp:755-927: candidates are features without a MapPoint; F12 is the 3x3 fundamental matrix row-major (the CV_32F cv::Mat's data), (ex, ey) the
    // projection of KF1's camera centre in KF2 (computed by the caller as the reference does at :765-772).  Returns the pair list.
    int SearchForTriangulation(const cv::Mat& desc1, const std::vector<cv::KeyPoint>& keysUn1, const std::vector<float>& uRight1,
                               const std::vector<uint8_t>& hasMapPoint1, const FeatVec& fv1, const cv::Mat& desc2,
                               const std::vector<cv::KeyPoint>& keysUn2, const std::vector<float>& uRight2,
                               const std::vector<uint8_t>& hasMapPoint2, const FeatVec& fv2, const float F12[9], float ex, float ey,
                               const std::vector<float>& scaleFactors2, const std::vector<float>& levelSigma2_2,
                               std::vector<std::pair<size_t, size_t> >& vMatchedPairs, bool bOnlyStereo, int device = 0) const
    {
        int n = 0;
        std::vector<int32_t> m12(desc1.rows, -1);
        orbm_check(orbm_search_for_triangulation(desc1.ptr(0), (const orbx_keypoint*)keysUn1.data(), uRight1.empty() ? nullptr : uRight1.data(), hasMapPoint1.data(),
                                                 desc1.rows, fv1.nodes.data(), fv1.offsets.data(), fv1.features.data(), (int)fv1.nodes.size(),
                                                 desc2.ptr(0), (const orbx_keypoint*)keysUn2.data(), uRight2.empty() ? nullptr : uRight2.data(), hasMapPoint2.data(),
                                                 desc2.rows, fv2.nodes.data(), fv2.offsets.data(), fv2.features.data(), (int)fv2.nodes.size(), F12,
                                                 ex, ey, scaleFactors2.data(), levelSigma2_2.data(), bOnlyStereo ? 1 : 0,
                                                 mbCheckOrientation ? 1 : 0, m12.data(), &n, device), "orbm_search_for_triangulation");
        vMatchedPairs.clear();
        vMatchedPairs.reserve(n);
        for (size_t i = 0; i < m12.size(); i++)           // :917-923
            if (m12[i] >= 0) vMatchedPairs.push_back(std::make_pair(i, (size_t)m12[i]));
        return n;
    }

    static const int TH_LOW = 50;         // src/ORBmatcher.cpp:37-39
    static const int TH_HIGH = 100;
    static const int HISTO_LENGTH = 30;

protected:
    float mfNNratio;
    bool mbCheckOrientation;
};

}  // namespace ORB_SLAM2
#endif
