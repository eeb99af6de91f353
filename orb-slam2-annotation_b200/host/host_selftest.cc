// host_selftest.cc — exercises the C++ host mirror (ORB_SLAM2::ORBextractor / ORBmatcher) exactly as Frame.cc does.
//   host_selftest                       no arguments: API/table checks; with no CUDA device it must fail loudly (exit 0)
//   host_selftest W H in.raw out.bin    extract in.raw (W x H, 8-bit) and write n, keypoints, descriptors, level sizes
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "ORBextractor.h"
#include "ORBmatcher.h"

using namespace ORB_SLAM2;

static int api_checks()
{
    ORBextractor ex(1000, 1.2f, 8, 20, 7);                       // TUM1.yaml ORB settings (Tracking.cc:141-149)
    if (ex.GetLevels() != 8 || ex.GetScaleFactor() != 1.2f) return 1;
    std::vector<float> sc = ex.GetScaleFactors(), isc = ex.GetInverseScaleFactors();
    std::vector<float> s2 = ex.GetScaleSigmaSquares(), is2 = ex.GetInverseScaleSigmaSquares();
    if (sc.size() != 8 || sc[0] != 1.0f || sc[1] != 1.2f || isc[1] != 1.0f / 1.2f || s2[1] != 1.2f * 1.2f || is2[2] != 1.0f / s2[2]) return 2;
    if (ORBmatcher::TH_LOW != 50 || ORBmatcher::TH_HIGH != 100 || ORBmatcher::HISTO_LENGTH != 30) return 3;
    std::vector<cv::KeyPoint> kps;
    cv::Mat desc, empty;
    ex(empty, cv::Mat(), kps, desc);                             // empty image: silent return
    if (!kps.empty()) return 4;
    cv::Mat img(480, 640, CV_8U);
    for (int y = 0; y < 480; y++) for (int x = 0; x < 640; x++) img.ptr(y)[x] = (unsigned char)((x * 7 + y * 13) ^ (x * y >> 3));
    if (orbb200_device_count() < 1) {
        try { ex(img, cv::Mat(), kps, desc); }
        catch (const OrbDeviceError& e) {
            if (e.code != ORBB_E_CUDA) return 5;
            printf("no CUDA device: operator() failed loudly (%s)\n", e.what());
            return 0;
        }
        return 6;                                                // a silent CPU result would be a bug
    }
    ex(img, cv::Mat(), kps, desc);
    if (kps.empty() || desc.rows != (int)kps.size() || desc.cols != 32) return 7;
    if ((int)ex.mvImagePyramid.size() != 8 || ex.mvImagePyramid[1].cols != 533 || ex.mvImagePyramid[7].rows != 134) return 8;
    if (memcmp(ex.mvImagePyramid[0].ptr(5), img.ptr(5), 640)) return 9;
    int d = ORBmatcher::DescriptorDistance(desc.row(0), desc.row(0));
    if (d != 0) return 10;
    printf("device path ok: %zu keypoints\n", kps.size());
    return 0;
}

int main(int argc, char** argv)
{
    if (argc < 5) { const int rc = api_checks(); if (rc) fprintf(stderr, "host_selftest: check %d failed\n", rc); return rc; }
    const int W = atoi(argv[1]), H = atoi(argv[2]);
    const int nf = argc > 5 ? atoi(argv[5]) : 1000;
    cv::Mat img(H, W, CV_8U);
    FILE* f = fopen(argv[3], "rb");
    if (!f || fread(img.ptr(0), 1, (size_t)W * H, f) != (size_t)W * H) { fprintf(stderr, "cannot read %s\n", argv[3]); return 2; }
    fclose(f);
    try {
        ORBextractor ex(nf, 1.2f, 8, 20, 7);
        std::vector<cv::KeyPoint> kps;
        cv::Mat desc;
        ex(img, cv::Mat(), kps, desc);
        FILE* o = fopen(argv[4], "wb");
        int n = (int)kps.size();
        fwrite(&n, 4, 1, o);
        fwrite(kps.data(), sizeof(cv::KeyPoint), n, o);
        for (int i = 0; i < n; i++) fwrite(desc.ptr(i), 1, 32, o);
        for (int l = 0; l < 8; l++) {
            int wh[2] = {ex.mvImagePyramid[l].cols, ex.mvImagePyramid[l].rows};
            fwrite(wh, 4, 2, o);
            for (int y = 0; y < wh[1]; y++) fwrite(ex.mvImagePyramid[l].ptr(y), 1, wh[0], o);
        }
        fclose(o);
    } catch (const std::exception& e) { fprintf(stderr, "host_selftest: %s\n", e.what()); return 3; }
    return 0;
}
