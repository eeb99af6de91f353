/* orbb200.h — C ABI of liborbb200.so: the B200-native (sm_100a CUDA) replacement for the ORB front-end hot
 * path of ORB-SLAM2.  The reference has no FFI; its boundary is the C++ class surface, so each entry point
 * cites the reference interface it replaces (paths relative to /root/reference/ORB-SLAM2/).
 *
 * Conventions: every function returns ORBB_OK (0) or a negative error code and never throws/aborts;
 * the caller allocates outputs and passes capacities; one orbx_handle owns one CUDA stream plus scratch,
 * calls on different handles are thread-safe, calls on one handle are serialised by the caller (the
 * reference's ORBextractor is likewise not re-entrant per instance: operator() mutates mvImagePyramid).
 * Pointers named *_host are host memory; pointers named *_dev are device memory on the handle's device.
 * There is NO CPU fallback: without a CUDA device every compute entry point returns ORBB_E_CUDA.
 */
#ifndef ORBB200_H
#define ORBB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    ORBB_OK = 0,
    ORBB_E_INVALID = -1,     /* bad argument (null pointer, nlevels out of range, iniTh < minTh, ...) */
    ORBB_E_CUDA = -2,        /* CUDA runtime error or no device; see orbb200_last_cuda_error() */
    ORBB_E_CAPACITY = -3,    /* an output or internal buffer was too small; *n_out holds the needed count */
    ORBB_E_BAD_SHAPE = -4,   /* image larger than the handle's max_w/max_h, or too small */
    ORBB_E_NOMEM = -5,
    ORBB_E_STATE = -6        /* call order violated (e.g. stereo match before both extractions) */
};

#define ORBB_MAX_LEVELS 16
#define ORBB_DESC_BYTES 32

/* Binary layout of cv::KeyPoint (28 bytes): what ORBextractor::operator() fills (ORBextractor.cpp:1084-1116). */
typedef struct orbx_keypoint {
    float x, y;        /* pt, level-0 pixel units (level coords * mvScaleFactor[octave]) */
    float size;        /* (float)(int)(31 * mvScaleFactor[octave])                        */
    float angle;       /* degrees in [0,360), IC_Angle / fastAtan2                        */
    float response;    /* FAST corner score                                               */
    int32_t octave;
    int32_t class_id;  /* always -1 */
} orbx_keypoint;

/* Mirrors ORBextractor::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
 * (include/ORBextractor.h:51-52, src/ORBextractor.cpp:412-472) plus the sizing a device handle needs. */
typedef struct orbx_config {
    int32_t nfeatures;
    float scale_factor;
    int32_t nlevels;         /* 1..ORBB_MAX_LEVELS */
    int32_t ini_th_fast;
    int32_t min_th_fast;     /* must be <= ini_th_fast */
    int32_t max_width;       /* largest image this handle will see */
    int32_t max_height;
    int32_t max_batch;       /* frames processed per orbx_extract_batch call (>=1) */
    int32_t device;          /* CUDA device ordinal */
    int32_t blur_legacy;     /* 0: OpenCV>=3.4 fixed-point kernel (graded); 1: OpenCV<=3.3 kernel {18,34,49,55,49,34,18} */
} orbx_config;

typedef struct orbx_handle orbx_handle;

/* -- extractor ------------------------------------------------------------------------------------- */

/* ORBextractor constructor (ORBextractor.cpp:412-472): scale tables, per-level quotas, pattern, umax. */
int orbx_create(const orbx_config* cfg, orbx_handle** out);
int orbx_destroy(orbx_handle* h);

/* Getters of ORBextractor.h:63-83 (GetScaleFactors, GetInverseScaleFactors, GetScaleSigmaSquares,
 * GetInverseScaleSigmaSquares) and the protected mnFeaturesPerLevel / umax tables (for tests).
 * Any output pointer may be NULL.  Arrays hold nlevels entries (umax: 16). */
int orbx_get_tables(const orbx_handle* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                    int32_t* features_per_level, int32_t* umax);

/* ORBextractor::operator()(image, mask [ignored], keypoints, descriptors) (ORBextractor.cpp:1053-1117),
 * host image in, host results out (includes H2D/D2H copies; this is the call Frame::ExtractORB makes,
 * Frame.cpp:259-265).  img_host: 8-bit single channel, stride_bytes between rows.
 * On success *n_out = number of keypoints (may exceed nfeatures slightly); if *n_out > capacity only the first
 * `capacity` are written and ORBB_E_CAPACITY is returned.  Empty image (w or h == 0) => *n_out = 0, ORBB_OK. */
int orbx_extract(orbx_handle* h, const uint8_t* img_host, int w, int h_, size_t stride_bytes,
                 orbx_keypoint* kps_host, uint8_t* desc_host, int capacity, int* n_out);

/* Same path with inputs already resident in HBM and results left in HBM: `nframes` (<= max_batch) images of
 * identical shape at imgs_dev + f*frame_stride_bytes.  Results for frame f: kps_dev + f*capacity keypoints,
 * desc_dev + f*capacity*32 bytes, counts_dev[f].  Asynchronous on the handle's stream. */
int orbx_extract_batch_device(orbx_handle* h, const uint8_t* imgs_dev, int nframes, int w, int h_,
                              size_t stride_bytes, size_t frame_stride_bytes,
                              orbx_keypoint* kps_dev, uint8_t* desc_dev, int32_t* counts_dev, int capacity);

/* The launch sequence of orbx_extract_batch_device captured once as a CUDA graph and replayed with one launch: for the caller of
 * Frame::ExtractORB (Frame.cpp:259-265), which extracts ONE frame at a time on the tracking thread's latency path, from fixed
 * device buffers.  orbx_graph_capture bakes the pointers, shape and capacity given into the handle's graph (replacing an earlier
 * one); orbx_graph_launch replays it on the handle's stream (asynchronous; orbx_sync as usual).  ORBB_E_STATE while profiling
 * is on or before a capture. */
int orbx_graph_capture(orbx_handle* h, const uint8_t* imgs_dev, int nframes, int w, int h_, size_t stride_bytes,
                       size_t frame_stride_bytes, orbx_keypoint* kps_dev, uint8_t* desc_dev, int32_t* counts_dev, int capacity);
int orbx_graph_launch(orbx_handle* h);

/* Host-buffer batch: pinned or pageable host frames in, host results out (one H2D copy per batch, one D2H).
 * counts_host[f] receives each frame's keypoint count. */
int orbx_extract_batch(orbx_handle* h, const uint8_t* imgs_host, int nframes, int w, int h_,
                       size_t stride_bytes, size_t frame_stride_bytes,
                       orbx_keypoint* kps_host, uint8_t* desc_host, int32_t* counts_host, int capacity);

/* Colour frames in front of the extractor: Tracking::GrabImageStereo / RGBD / Monocular convert with
 * cv::cvtColor(CV_RGB2GRAY | CV_BGR2GRAY | CV_RGBA2GRAY | CV_BGRA2GRAY) before calling ORBextractor::operator()
 * (src/Tracking.cpp:202-227, :233-250, :264-281).  channels = 3 or 4 (alpha ignored), rgb_order = mbRGB (1: R first,
 * 0: B first); arithmetic of OpenCV 4.x: (R*9798 + G*19235 + B*3735 + 2^14) >> 15.
 * orbx_extract_batch_color = the conversion on the device followed by orbx_extract_batch (stride / fstride in bytes of the
 * colour frames; everything else as orbx_extract_batch).  orbx_cvt_gray_device converts device frames on a cudaStream_t. */
int orbx_extract_batch_color(orbx_handle* h, const uint8_t* imgs_host, int nframes, int w, int h_,
                             size_t stride_bytes, size_t frame_stride_bytes, int channels, int rgb_order,
                             orbx_keypoint* kps_host, uint8_t* desc_host, int32_t* counts_host, int capacity);
int orbx_cvt_gray_device(const uint8_t* src_dev, int nframes, int w, int h_, size_t src_stride, size_t src_frame_stride,
                         int channels, int rgb_order, uint8_t* dst_dev, size_t dst_stride, size_t dst_frame_stride, void* stream);

/* Stereo rectification in front of the extractor: the stereo examples call
 * cv::remap(imLeft, imLeftRect, M1l, M2l, cv::INTER_LINEAR) with the CV_32F maps of cv::initUndistortRectifyMap before
 * SLAM.TrackStereo (Examples/Stereo/stereo_euroc.cc:174-185).  orbx_set_rectification stores one camera's maps (dst_w x dst_h
 * floats each, giving for every rectified pixel its source position in the src_w x src_h raw frame) in the handle, converted to
 * cv::remap's fixed point; NULL maps clear it.  orbx_extract_batch_rectified = cv::remap on the device (INTER_LINEAR,
 * BORDER_CONSTANT 0, arithmetic of OpenCV 4.x) followed by orbx_extract_batch on the rectified frames; raw frames are
 * src_w x src_h bytes with the given strides.  orbx_remap_device rectifies device frames and synchronises the stream. */
int orbx_set_rectification(orbx_handle* h, const float* mapx_host, const float* mapy_host, int dst_w, int dst_h, int src_w, int src_h);
int orbx_extract_batch_rectified(orbx_handle* h, const uint8_t* raw_host, int nframes, size_t stride_bytes, size_t frame_stride_bytes,
                                 orbx_keypoint* kps_host, uint8_t* desc_host, int32_t* counts_host, int capacity);
int orbx_remap_device(const uint8_t* src_dev, int nframes, int src_w, int src_h, size_t src_stride, size_t src_frame_stride,
                      const float* mapx_host, const float* mapy_host, int dst_w, int dst_h,
                      uint8_t* dst_dev, size_t dst_stride, size_t dst_frame_stride, void* stream);

/* Blocks until the handle's stream is idle; returns the first asynchronous error, and ORBB_E_CAPACITY if an
 * internal candidate buffer overflowed during the last extraction. */
int orbx_sync(orbx_handle* h);
/* Measurement support: with profiling on, every extraction records CUDA events around its three kernel classes on the
 * handle's stream and accumulates their durations: ms3[0] the per-level resize/blur/FAST kernels, ms3[1] the octree
 * kernel, ms3[2] the orientation+descriptor kernel; *calls = extractions accumulated.  (Adds one host sync per call:
 * never leave it on inside a throughput measurement.) */
int orbx_set_profiling(orbx_handle* h, int on);
int orbx_get_profile(orbx_handle* h, double* ms3, int* calls);
/* cudaStream_t of the handle (for event timing by the caller). */
void* orbx_stream(orbx_handle* h);

/* Backs the public member ORBextractor::mvImagePyramid (ORBextractor.h:85; read by Frame.cpp:547,657-676):
 * copies level `level` of frame `frame` of the last extraction to dst_host (tight rows of *w bytes when
 * dst_stride == 0).  dst_host may be NULL to query the size only. */
int orbx_pyramid_level(orbx_handle* h, int frame, int level, uint8_t* dst_host, size_t dst_stride, int* w, int* h_);

/* Parity-test access to intermediate stages of the last extraction (frame `frame`, level `level`). */
enum {
    ORBX_STAGE_BLUR = 1,      /* uint8 blurred level (GaussianBlur 7x7 s=2, ORBextractor.cpp:1097-1098), tight rows  */
    ORBX_STAGE_FASTKPS = 2,   /* int32 triples (x, y, response) in level coords relative to minBorder: the list
                                 handed to DistributeOctTree (ORBextractor.cpp:829-834); order unspecified        */
    ORBX_STAGE_LEVELKPS = 3   /* orbx_keypoint records after octree + angle in level coordinates (:847-862), in
                                 reference order                                                               */
};
int orbx_debug_stage(orbx_handle* h, int frame, int level, int stage, void* dst_host, size_t dst_bytes, size_t* bytes_out);

/* -- matcher ---------------------------------------------------------------------------------------- */

/* ORBmatcher::DescriptorDistance (ORBmatcher.cpp:1838-1854; twin DBoW2/FORB.cpp:82-102) for n pairs
 * a[i], b[i] of 32-byte descriptors (host pointers). */
int orbm_distance_pairs(const uint8_t* a_host, const uint8_t* b_host, int n, int32_t* out_host, int device);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cpp:302-380), batched over map points.
 * Map point p owns the descriptor rows [offsets[p], offsets[p+1]) of `desc` (32 B each): the rows of its observations, in
 * the order the caller walks mObservations.  best_row[p] = index, relative to offsets[p], of the row whose median Hamming
 * distance to the point's rows (itself included, median = sorted[int(0.5*(N-1))], :361-365) is smallest, first row on
 * ties (:368); -1 for a point without rows (the reference returns early, :316-317, :330-331).  best_median may be NULL.
 * The _device form takes device pointers (desc 16-byte aligned) and a cudaStream_t and does not synchronise. */
int orbm_distinctive_descriptors(const uint8_t* desc_host, const int64_t* offsets_host, int npoints,
                                 int32_t* best_row_host, int32_t* best_median_host, int device);
int orbm_distinctive_descriptors_device(const uint8_t* desc_dev, const int64_t* offsets_dev, int npoints,
                                        int32_t* best_row_dev, int32_t* best_median_dev, void* stream);

/* Frame::UndistortKeyPoints (Frame.cpp:462-496): mvKeysUn = mvKeys with (x, y) replaced by cv::undistortPoints(pts, mK, mDistCoef,
 * cv::Mat(), mK); a plain copy when mDistCoef(0) == 0.  K4 = {fx, fy, cx, cy} (mK), dist = mDistCoef with ndist = 4, 5, 8 or 12
 * coefficients (k1 k2 p1 p2 [k3 [k4 k5 k6 [s1..s4]]]).  OpenCV 4.x arithmetic (double, five iterations), bit-equal to cv2 4.13
 * (tests/golden/cv2_undistort.npz).  The _device form works on device arrays of any length (a whole batch of frames that share one
 * calibration) and is queued on `stream`; in and out may be the same array.
 * orbm_image_bounds: Frame::ComputeImageBounds (Frame.cpp:498-536) -> {mnMinX, mnMinY, mnMaxX, mnMaxY}. */
int orbm_undistort_keypoints(const orbx_keypoint* kps_host, int n, const float* K4, const float* dist, int ndist,
                             orbx_keypoint* kps_un_host, int device);
int orbm_undistort_keypoints_device(const orbx_keypoint* kps_dev, int n, const float* K4, const float* dist, int ndist,
                                    orbx_keypoint* kps_un_dev, int device, void* stream);
int orbm_image_bounds(int cols, int rows, const float* K4, const float* dist, int ndist, float* bounds4, int device);

/* A Frame's matching state on the device: undistorted keypoints (mvKeysUn), descriptors (mDescriptors) and the
 * 64x48 bucket grid of Frame::AssignFeaturesToGrid / PosInGrid (Frame.cpp:241-257, :434-444), built on device
 * in insertion order.  Bounds are Frame::mnMinX/mnMinY/mnMaxX/mnMaxY. */
typedef struct orbm_frame orbm_frame;
int orbm_frame_create(orbm_frame** out, const orbx_keypoint* kps_un_host, const uint8_t* desc_host, int n,
                      float min_x, float min_y, float max_x, float max_y, int device);
/* The same from device arrays (e.g. the outputs of orbx_extract_batch_device after undistortion): copies and grid build are
 * queued on `stream`, nothing visits the host. */
int orbm_frame_create_device(orbm_frame** out, const orbx_keypoint* kps_un_dev, const uint8_t* desc_dev, int n,
                             float min_x, float min_y, float max_x, float max_y, int device, void* stream);
int orbm_frame_destroy(orbm_frame* f);

/* One projected query: what every SearchByProjection variant hands to GetFeaturesInArea + DescriptorDistance
 * (ORBmatcher.cpp:63-155, :352-472, :1506-1659, :1661-1789; SearchForInitialization :474-590). */
typedef struct orbm_query {
    float u, v;             /* window centre                                                        */
    float r;                /* radius, already multiplied by the scale factor                       */
    int32_t min_level;      /* GetFeaturesInArea(x, y, r, minLevel, maxLevel) (Frame.cpp:379)        */
    int32_t max_level;
    uint8_t desc[ORBB_DESC_BYTES];
} orbm_query;

/* Frame::GetFeaturesInArea (Frame.cpp:379-432) + DescriptorDistance for nq queries at once.  Output is CSR:
 * cand_offsets[nq+1], cand_idx / cand_dist in EXACTLY the reference's enumeration order (ix, iy, insertion),
 * so the caller's order-dependent resolution loops (already-matched skips, ratio tests, rotation histogram)
 * run unchanged on the host.  If the total exceeds cap, ORBB_E_CAPACITY is returned with cand_offsets filled. */
int orbm_window_search(const orbm_frame* f, const orbm_query* q_host, int nq,
                       int32_t* cand_offsets_host, int32_t* cand_idx_host, int32_t* cand_dist_host, int cap);

/* A projected query plus the per-query state the reference's resolution loops read. */
typedef struct orbm_proj_query {
    orbm_query q;
    float u_right;          /* a10: pMP->mTrackProjXR; a11: u - mbf*invzc (ORBmatcher.cpp:116, :1597)        */
    float max_err_right;    /* candidate skipped if mvuRight[idx] > 0 && |u_right - mvuRight[idx]| > this     */
    int32_t id;             /* caller's identifier of the map point (stored into frame_mp on a match)       */
    int32_t has_obs;        /* pMP->Observations() > 0                                                      */
    float angle;            /* LastFrame.mvKeysUn[i].angle, for the rotation histogram (a11)                */
} orbm_proj_query;

/* ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12, vector<pair<size_t,size_t>>& vMatchedPairs,
 * const bool bOnlyStereo) (ORBmatcher.cpp:755-927).  kps_un* = mvKeysUn, uright* = mvuRight (may be NULL for monocular KeyFrames
 * unless only_stereo), has_mp*[i] != 0 iff GetMapPoint(i) is set; F12 = the fundamental matrix, 9 floats row-major; (ex, ey) = the
 * epipole of KeyFrame 1 in KeyFrame 2 as computed at :763-769; scale_factors2 = pKF2->mvScaleFactors, level_sigma2_2 =
 * pKF2->mvLevelSigma2.  match12[n1] = vMatches12 (KeyFrame-2 feature or -1): vMatchedPairs is its list of (i, match12[i]). */
int orbm_search_for_triangulation(const uint8_t* desc1, const orbx_keypoint* kps_un1, const float* uright1, const uint8_t* has_mp1, int n1,
                                  const int32_t* fv1_nodes, const int32_t* fv1_off, const int32_t* fv1_feat, int nfv1,
                                  const uint8_t* desc2, const orbx_keypoint* kps_un2, const float* uright2, const uint8_t* has_mp2, int n2,
                                  const int32_t* fv2_nodes, const int32_t* fv2_off, const int32_t* fv2_feat, int nfv2,
                                  const float* F12, float ex, float ey, const float* scale_factors2, const float* level_sigma2_2,
                                  int only_stereo, int check_orientation, int32_t* match12, int* nmatches, int device);

/* the relocalisation, loop-closing and fusion variants.  The caller projects (pose algebra,
 * PredictScale, frustum tests) exactly as the reference does and hands over one orbm_proj_query per surviving MapPoint:
 * q.u, q.v, q.r = th * mvScaleFactors[nPredictedLevel], q.min_level / q.max_level = the octave window of the variant.
 *
 * orbm_search_by_projection_reloc: SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound,
 *   th, ORBdist) (ORBmatcher.cpp:1661-1789): level window [p-1, p+1]; frame_mp[i] >= 0 marks Frame features that already have a
 *   MapPoint (:1727-1728) and receives pq[].id on a match; max_dist = ORBdist; angle = pKF->mvKeysUn[i].angle.
 * orbm_search_by_projection_loop: SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints,
 *   vector<MapPoint*>& vpMatched, th) (:352-472): level window [p-1, p]; matched[i] >= 0 = vpMatched[i] set, receives pq[].id.
 * orbm_fuse_best: the search of Fuse(KeyFrame*, const vector<MapPoint*>&, th) (:962-1115: pass mvuRight, mvInvLevelSigma2 and
 *   q.u_right = ur for the chi-square gate, max_dist = TH_LOW), of Fuse(KeyFrame*, cv::Mat Scw, ...) (:1119-1251: gate off =
 *   inv_level_sigma2 NULL, TH_LOW) and of each direction of SearchBySim3 (:1253-1500: gate off, TH_HIGH).  best_idx[i] = the
 *   KeyFrame feature the reference selects for point i, or -1; Replace / AddObservation / the Sim3 cross-check stay with the caller. */
int orbm_search_by_projection_reloc(const orbm_frame* f, const orbm_proj_query* q_host, int nq, int max_dist, int check_orientation,
                                    int32_t* frame_mp, int* nmatches);
int orbm_search_by_projection_loop(const orbm_frame* f, const orbm_proj_query* q_host, int nq, int32_t* matched, int* nmatches);
/* Device-resident form of the two calls above (they are this call plus the copies): queries and the slot array (frame_mp / vpMatched as
 * ids, -1 = empty) live on the device, window search and the order-dependent resolve ("a slot that holds a MapPoint is skipped", so every
 * accepted match blocks its slot for the later queries: evaluated as the parallel fixed-point iteration of the a10 / a11 resolve) run
 * there, queued on `stream` without host synchronisation.  Relocalisation: max_dist = ORBdist, check_orientation = mbCheckOrientation;
 * loop closing: max_dist = TH_LOW (50), check_orientation = 0.  *status_dev / cand_capacity as in orbm_search_by_projection_points_device. */
int orbm_search_by_projection_reloc_device(const orbm_frame* f, const orbm_proj_query* q_dev, int nq, int max_dist, int check_orientation,
                                           int32_t* frame_mp_dev, int32_t* nmatches_dev, int32_t* status_dev, int cand_capacity, void* stream);
int orbm_fuse_best(const orbm_frame* f, const orbm_proj_query* q_host, int nq, const float* mvuRight_host,
                   const float* inv_level_sigma2, int max_dist, int32_t* best_idx, int32_t* best_dist);

/* ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches) (ORBmatcher.cpp:205-348).
 * desc_kf / desc_f: mDescriptors of the KeyFrame (n_kf rows) and of the Frame (n_f rows); angle_kf = pKF->mvKeysUn[i].angle,
 * angle_f = F.mvKeys[i].angle (read only when check_ori); kf_has_mp[i] != 0 iff vpMapPointsKF[i] exists and is not bad
 * (:233-238).  fv_*: the two FeatureVectors as orbv_compute_bow returns them (node ids ascending, CSR offsets, feature
 * indices).  Output match_f_to_kf[n_f]: the KeyFrame feature whose MapPoint the reference would store in
 * vpMapPointMatches[i], or -1; *nmatches = the function's return value. */
int orbm_search_by_bow(const uint8_t* desc_kf, const float* angle_kf, const uint8_t* kf_has_mp, int n_kf,
                       const int32_t* fv_kf_nodes, const int32_t* fv_kf_off, const int32_t* fv_kf_feat, int nfv_kf,
                       const uint8_t* desc_f, const float* angle_f, int n_f,
                       const int32_t* fv_f_nodes, const int32_t* fv_f_off, const int32_t* fv_f_feat, int nfv_f,
                       float nnratio, int check_orientation, int32_t* match_f_to_kf, int* nmatches, int device);

/* ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12) (ORBmatcher.cpp:604-743), the loop-closing
 * variant: has_mp1 / has_mp2 flag features whose MapPoint exists and is not bad, angle* = mvKeysUn[i].angle.  Output
 * match12[n1]: the KeyFrame-2 feature whose MapPoint the reference stores in vpMatches12[i], or -1. */
int orbm_search_by_bow_kf(const uint8_t* desc1, const float* angle1, const uint8_t* has_mp1, int n1,
                          const int32_t* fv1_nodes, const int32_t* fv1_off, const int32_t* fv1_feat, int nfv1,
                          const uint8_t* desc2, const float* angle2, const uint8_t* has_mp2, int n2,
                          const int32_t* fv2_nodes, const int32_t* fv2_off, const int32_t* fv2_feat, int nfv2,
                          float nnratio, int check_orientation, int32_t* match12, int* nmatches, int device);

/* ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th) (ORBmatcher.cpp:63-155): window candidates and
 * distances on the device, then the reference's order-dependent top-2 / level-aware ratio / already-matched loop, also
 * on the device (see the _device forms below; this host-pointer form uploads, runs them and reads the state back).  frame_mp[n] (in/out) is F.mvpMapPoints flattened to ids (-1 = none), frame_mp_obs[n] the matched point's
 * Observations()>0.  mvuRight_host may be NULL (monocular).  Queries are the visible, non-bad map points in order. */
int orbm_search_by_projection_points(const orbm_frame* f, const float* mvuRight_host, const orbm_proj_query* q_host, int nq,
                                     float nnratio, int32_t* frame_mp, int32_t* frame_mp_obs, int* nmatches);
/* ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th, bMono) (ORBmatcher.cpp:1506-1659) after the
 * projection step: top-1 <= TH_HIGH, already-matched skips, rotation histogram (ComputeThreeMaxima :1792-1833). */
int orbm_search_by_projection_frame(const orbm_frame* f, const float* mvuRight_host, const orbm_proj_query* q_host, int nq,
                                    int check_orientation, int32_t* frame_mp, int32_t* frame_mp_obs, int* nmatches);
/* Device-resident forms of the two calls above: queries, mvuRight (or NULL), the Frame's match state and the two
 * result scalars all live on the device; window search AND the order-dependent resolve run there (the reference's sequential
 * "skip slots that already hold an observed MapPoint" rule is evaluated as a parallel fixed-point iteration that reproduces the
 * sequential result exactly), queued on `stream` with no host synchronisation.  *status_dev = ORBB_OK, or ORBB_E_CAPACITY when the
 * windows held more than cand_capacity candidates in total (0 = 64*nq + 1024), in which case the state is left untouched.
 * Calls that share one orbm_frame share its workspace: the caller serialises them (they are the Tracking thread's calls). */
int orbm_search_by_projection_points_device(const orbm_frame* f, const float* mvuRight_dev, const orbm_proj_query* q_dev, int nq,
                                            float nnratio, int32_t* frame_mp_dev, int32_t* frame_mp_obs_dev, int32_t* nmatches_dev,
                                            int32_t* status_dev, int cand_capacity, void* stream);
int orbm_search_by_projection_frame_device(const orbm_frame* f, const float* mvuRight_dev, const orbm_proj_query* q_dev, int nq,
                                           int check_orientation, int32_t* frame_mp_dev, int32_t* frame_mp_obs_dev, int32_t* nmatches_dev,
                                           int32_t* status_dev, int cand_capacity, void* stream);
/* ORBmatcher::SearchForInitialization (ORBmatcher.cpp:474-590): level-0 keypoints of f1 against f2 in a window around
 * prev_matched_xy (in/out, 2*n1 floats); matches12_out[n1] = index in f2 or -1. */
int orbm_search_for_initialization(const orbm_frame* f1, const orbm_frame* f2, float* prev_matched_xy, int32_t* matches12_out,
                                   int window_size, float nnratio, int check_orientation, int* nmatches);

/* Frame::ComputeStereoMatches (Frame.cpp:540-748) on the device, reading both extractors' pyramids
 * (mpORBextractorLeft/Right->mvImagePyramid) in place.  left/right must hold their last extraction (frame 0);
 * kps/desc are the host vectors those extractions returned.  mb is the member value at call time (0 in the
 * reference's stereo constructor => no max-disparity cut).  Outputs mvuRight / mvDepth (n_left floats, -1 = none). */
int orbm_stereo_match(orbx_handle* left, orbx_handle* right,
                      const orbx_keypoint* kps_left_host, const uint8_t* desc_left_host, int n_left,
                      const orbx_keypoint* kps_right_host, const uint8_t* desc_right_host, int n_right,
                      float mbf, float mb, float* u_right_host, float* depth_host);

/* Frame::Frame, the stereo constructor (Frame.cpp:61-115), as ONE call: ExtractORB(0, imLeft) || ExtractORB(1, imRight) (:84-87, the two
 * threads become one two-frame batch of `h`, which therefore needs max_batch >= 2 and plays both mpORBextractorLeft and ...Right),
 * UndistortKeyPoints for rectified input (mDistCoef[0] == 0 => mvKeysUn = mvKeys, :467-472: what the reference's stereo examples feed),
 * ComputeStereoMatches (:540-748) and AssignFeaturesToGrid (:241-257) on the results where they lie in device memory.  Nothing is uploaded
 * twice and the host synchronises twice (counts known; everything done).  Results are bit-equal to orbx_extract (left), orbx_extract
 * (right), orbm_stereo_match, orbm_frame_create called one after the other (tests/test_gpu_match.py::test_stereo_frame_create).
 * Out: *out (for the Search* calls; orbm_frame_destroy), mvKeys / mDescriptors (kps_left, desc_left, *n_left), mvKeysRight /
 * mDescriptorsRight (kps_right, desc_right, *n_right; both may be NULL), mvuRight / mvDepth (n_left floats, -1 = none).  `capacity` rows
 * per host array; ORBB_E_CAPACITY (counts still set) when an image has more keypoints.  min_x..max_y: mnMinX.. (orbm_image_bounds). */
int orbm_stereo_frame_create(orbx_handle* h, const uint8_t* left_host, const uint8_t* right_host, int w, int h_, size_t stride_bytes,
                             float mbf, float mb, float min_x, float min_y, float max_x, float max_y, orbm_frame** out,
                             orbx_keypoint* kps_left_host, uint8_t* desc_left_host, int* n_left,
                             orbx_keypoint* kps_right_host, uint8_t* desc_right_host, int* n_right, int capacity,
                             float* u_right_host, float* depth_host);
/* Frame::Frame, the monocular constructor (Frame.cpp:171-224; the RGB-D constructor :117-169 differs by its depth lookup only), as ONE call:
 * ExtractORB(0, imGray) (:186), UndistortKeyPoints (:462-496; K4 / dist may be NULL, and dist[0] == 0 means mvKeysUn = mvKeys as at :465-469)
 * and AssignFeaturesToGrid (:241-257) on the keypoints where the extraction left them.  Out: *out, mvKeys (kps), mvKeysUn (kps_un),
 * mDescriptors (desc), *n; `capacity` rows per array, ORBB_E_CAPACITY (*n still set) when the image has more keypoints.  Bit-equal to
 * orbx_extract + orbm_undistort_keypoints + orbm_frame_create (tests/test_gpu_match.py::test_mono_frame_create). */
int orbm_mono_frame_create(orbx_handle* h, const uint8_t* img_host, int w, int h_, size_t stride_bytes,
                           const float* K4, const float* dist, int ndist, float min_x, float min_y, float max_x, float max_y,
                           orbm_frame** out, orbx_keypoint* kps_host, orbx_keypoint* kps_un_host, uint8_t* desc_host, int* n, int capacity);
/* Pass as mvuRight_host to orbm_search_by_projection_points / _frame on a frame built by orbm_stereo_frame_create: the search uses the
 * mvuRight that constructor left on the device instead of uploading the host copy (ORBB_E_STATE on any other frame). */
#define ORBM_URIGHT_OF_FRAME ((const float*)(uintptr_t)1)

/* Brute-force top-2 over a descriptor database shard with the reference's top-2 update rule
 * (SearchByBoW, ORBmatcher.cpp:249-275: strict '<', first index wins, second may equal best).
 * A database object holds one shard: rows [row_offset, row_offset + nrows) of the global database. */
typedef struct orbm_db orbm_db;
int orbm_db_create(orbm_db** out, int64_t capacity_rows, int64_t row_offset, int device);
int orbm_db_append(orbm_db* db, const uint8_t* rows_host, int64_t nrows);
int orbm_db_append_device(orbm_db* db, const uint8_t* rows_dev, int64_t nrows);
int orbm_db_destroy(orbm_db* db);
int64_t orbm_db_rows(const orbm_db* db);

/* Host queries in, host results out: d1[i] best distance, idx1[i] GLOBAL row index of the best (first on ties),
 * d2[i] second-best distance (256 when the shard has < 2 rows; d1 = 256, idx1 = -1 when it has none).
 * A search uses (and may grow) scratch owned by the database object, so the handle is not const and calls on ONE orbm_db must be
 * serialised by the caller, also across streams; concurrent searchers each create their own orbm_db (rows can be shared by
 * appending from the same device buffer). */
int orbm_bf_top2(orbm_db* db, const uint8_t* q_host, int nq, int32_t* d1_host, int64_t* idx1_host, int32_t* d2_host);
/* Device queries / device results, asynchronous on `stream` (cudaStream_t, may be NULL). */
int orbm_bf_top2_device(orbm_db* db, const uint8_t* q_dev, int nq, int32_t* d1_dev, int64_t* idx1_dev,
                        int32_t* d2_dev, void* stream);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  The same scan and chunk merge as orbm_bf_top2_device over rows
 * the CALLER holds on the device (rows_dev: nrows x 32 B, 16-byte aligned; global row index = row_offset + i), with the per-chunk
 * partial keys in a file-scope static device array of liborbb200 (4 MiB, held by every GPU that loads the library until exit)
 * instead of scratch owned by an orbm_db.  Calls no cudaMalloc / cudaFree and no allocating library routine.
 * Limits: nq <= ORBM_NOALLOC_MAX_QUERIES; the scan is split into at most ORBM_NOALLOC_MAX_KEYS / nq chunks of at most 2^23 rows, so
 * nrows <= (ORBM_NOALLOC_MAX_KEYS / nq) * 2^23; beyond either limit ORBB_E_CAPACITY is returned (text in orbb200_last_cuda_error)
 * and nothing is launched.  Every call on one device shares the array: calls that overlap on the GPU (two streams, two threads)
 * race; serialise them.  Asynchronous on `stream`; outputs are device pointers as in orbm_bf_top2_device. */
#define ORBM_NOALLOC_MAX_QUERIES 4096
#define ORBM_NOALLOC_MAX_KEYS (1 << 19)
int orbm_bf_top2_noalloc(const uint8_t* rows_dev, int64_t nrows, int64_t row_offset, const uint8_t* q_dev, int nq,
                         int32_t* d1_dev, int64_t* idx1_dev, int32_t* d2_dev, int device, void* stream);
/* Merge of G per-shard results laid out [g][nq] (shards ordered by ascending row range) into the global top-2
 *  — the step that follows the NCCL all-gather.  Device pointers, asynchronous on `stream`. */
int orbm_merge_top2_device(const int32_t* d1_all, const int64_t* idx1_all, const int32_t* d2_all, int nshards, int nq,
                           int32_t* d1_dev, int64_t* idx1_dev, int32_t* d2_dev, void* stream);
/* Ratio test of the reference (ORBmatcher.cpp:278-305): accept iff d1 <= th_low && (float)d1 < ratio*(float)d2.
 * Host arrays; accepted[i] = 0/1. */
int orbm_ratio_test(const int32_t* d1, const int32_t* d2, int nq, int th_low, float ratio, uint8_t* accepted);

/* -- vocabulary (Frame::ComputeBoW) ------------------------------------------------------- */

/* Device image of a DBoW2 vocabulary tree (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:312-336 Node, :1362-1448
 * loadFromTextFile).  Nodes are given in the loader's order: node 0 is the root, node i >= 1 names parent[i] < i (children are
 * appended to their parent in this order), is_leaf[i] = the file's nIsLeaf flag (leaves get consecutive word ids, :1432-1439),
 * desc = n_nodes x 32 B (row 0 unused), weight = the file's weight column (idf).  weighting: 0 TF_IDF, 1 TF, 2 IDF, 3 BINARY
 * (BowVector.h); norm: what the scoring object's mustNormalize() asks for, 0 none, 1 L1, 2 L2 (ScoringObject.h; ORBvoc.txt is
 * "10 6 0 0": TF_IDF + L1_NORM -> weighting 0, norm 1).  A leaf flag that disagrees with "has no children" is rejected. */
typedef struct orbv_vocab orbv_vocab;
int orbv_vocab_create(orbv_vocab** out, int k, int L, int n_nodes, const int32_t* parent, const uint8_t* is_leaf,
                      const uint8_t* desc, const double* weight, int weighting, int norm, int device);
int orbv_vocab_destroy(orbv_vocab* v);
int orbv_vocab_words(const orbv_vocab* v);

/* TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup) (:1242-1285) for n descriptors: walk down the tree,
 * at each level the first child with the smallest FORB::distance; node_id = the node reached at level L - levelsup (0 = root
 * when that level is <= 0, and also when the leaf is reached above that level, where the reference leaves *nid unset).
 * node_id may be NULL.  The _device form takes device pointers (desc 16-byte aligned) and a cudaStream_t, no synchronisation. */
int orbv_transform(const orbv_vocab* v, const uint8_t* desc_host, int n, int levelsup,
                   int32_t* word_id, double* weight, int32_t* node_id);
int orbv_transform_device(const orbv_vocab* v, const uint8_t* desc_dev, int n, int levelsup,
                          int32_t* word_id_dev, double* weight_dev, int32_t* node_id_dev, void* stream);

/* transform(features, BowVector&, FeatureVector&, levelsup) (:1151-1218) = Frame::ComputeBoW (src/Frame.cpp:452-459, levelsup 4).
 * BowVector: n_bow (word id, value) pairs, ids ascending (std::map order), values accumulated and normalised as the reference
 * does.  FeatureVector: n_fv node ids ascending with CSR offsets into fv_features (feature indices in insertion order).
 * Capacities: bow_ids/bow_vals/fv_nodes/fv_features n entries, fv_offsets n + 1. */
int orbv_compute_bow(const orbv_vocab* v, const uint8_t* desc_host, int n, int levelsup,
                     int32_t* bow_ids, double* bow_vals, int32_t* n_bow,
                     int32_t* fv_nodes, int32_t* fv_offsets, int32_t* fv_features, int32_t* n_fv);

/* -- misc ------------------------------------------------------------------------------------------- */
const char* orbb200_strerror(int code);
const char* orbb200_last_cuda_error(void);
int orbb200_device_count(void);
/* Measured throughput (1e9 POPC.32 per second) of a register-only __popc loop on `device`: the roofline
 * denominator for the matching kernels. */
int orbb200_measure_popc_peak(int device, double* gpopc_per_s);
/* Measured throughput (int8 TOPS, 2 x multiply-adds) of a register-only mma.sync.m16n8k32 (IMMA) loop on `device`: the
 * tensor-pipe denominator of the brute-force top-2 kernel, where one 256-bit comparison is 512 int8 operations. */
int orbb200_measure_imma_peak(int device, double* int8_tops);
/* Measured throughput (int8 TOPS) of tcgen05.mma kind::i8 128x256x32 issued back to back from shared-memory operands into TMEM
 * accumulators, one CTA per SM, nothing else running: the measured tensor-pipe denominator of bf_top2_tcn_kernel's roofline. */
int orbb200_measure_tcgen05_i8_peak(int device, double* int8_tops);
/* Number of kernel launches issued by this library since load (bench.py's gpu_launches claim). */
uint64_t orbb200_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* ORBB200_H */
