// common.cuh — shared helpers of liborbb200 (sm_100a only; no CPU fallback anywhere in this library).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <atomic>
#include <mutex>
#include <new>
#include <vector>
#include "../../include/orbb200.h"

namespace orbb {

extern std::atomic<uint64_t> g_launches;     // kernels launched by this library (orbb200_launch_count)
void set_cuda_error(cudaError_t e, const char* what);

#define ORBB_CUDA(call)                                              \
    do {                                                             \
        cudaError_t _e = (call);                                     \
        if (_e != cudaSuccess) {                                     \
            ::orbb::set_cuda_error(_e, #call);                       \
            return ORBB_E_CUDA;                                      \
        }                                                            \
    } while (0)

#define ORBB_LAUNCHED() (::orbb::g_launches.fetch_add(1, std::memory_order_relaxed))

// Scratch allocator of the matcher / vocabulary entry points: device blocks are cached per device in power-of-two size classes, so the
// per-frame calls (orbm_frame_create, orbm_window_search, the Search* functions, orbm_stereo_match ...) stop calling cudaMalloc /
// cudaFree after their first use - those calls serialise the device and cost ~100 us per frame on the tracking thread.
// A block is handed back only when no queued work can still touch it (the host-pointer entry points synchronise before they return;
// orbm_frame_destroy synchronises the streams the frame was used on).  Blocks above 64 MiB and anything beyond 1 GiB cached per
// device go straight to cudaMalloc / cudaFree.  Defined in extract.cu.
void* scratch_alloc(size_t bytes, size_t* granted);      // nullptr on failure (the CUDA error is recorded)
void scratch_free(void* p, size_t granted);
void scratch_trim(void);                                 // cudaFree everything cached on the current device

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Device-side pyramid geometry of one handle (levels of all frames share it).
struct LevelGeom {
    int w, h;            // level size: cvRound(W * mvInvScaleFactor[L]) (ORBextractor.cpp:1128)
    int pitch;           // row pitch in bytes inside the pyramid / blur buffers (multiple of 128)
    size_t off;          // byte offset of this level inside one frame's pyramid (and blur) block
    int nCols, nRows, wCell, hCell;   // FAST cell grid (ORBextractor.cpp:781-795); nCols==0 => no FAST on this level
    unsigned magicW, magicH;          // ceil(2^32 / wCell), ceil(2^32 / hCell): x / wCell == __umulhi(x, magicW) for x < 2^16
    int nIni;            // octree roots (ORBextractor.cpp:545)
    float hX;            // root width as float
    int quota;           // mnFeaturesPerLevel[L]
    int rec_cap;         // capacity of the raw FAST record list of this level
    size_t rec_off;      // element offset of this level's record list inside one frame's record block
    int out_cap;         // capacity of the per-level octree output
    size_t out_off;      // element offset of the per-level octree output inside one frame's block
    int node_cap;        // octree node pool capacity
    float scale;         // mvScaleFactor[L]
    float size;          // (float)(int)(31 * scale)
    // resize tables for L -> L+1 (unused on the last level)
    int xt_off, yt_off;  // offsets into the coefficient table (int2 {s0, a0 | a1 << 16})
    int dxs_off, dys_off;// offsets into the per-tile first-destination-index tables
};

}  // namespace orbb

// What orbx__extract_resident (extract.cu) hands to orbm_stereo_frame_create (match.cu): where one batch's results and pyramids live.
struct orbx_resident_view {
    orbx_keypoint* kps_dev; uint8_t* desc_dev;           // [frame][devcap] rows
    const orbx_keypoint* kps_pinned; const uint8_t* desc_pinned;   // the handle's pinned copies, same layout (valid after s_out is drained)
    int devcap;
    const orbb::LevelGeom *geom_host, *geom_dev; int nlevels;
    const uint8_t* src0; long long src0_pitch, src0_fstride;       // level 0 of frame 0 and its strides
    const uint8_t* pyr; long long pyr_fstride;                     // levels >= 1 at geom[L].off inside each frame's block
    const float *scale, *inv_scale;
    int device; void *stream, *s_out, *s_aux;                      // s_aux: an idle stream of the handle for work beside `stream`
};

namespace orbb {

// ORBmatcher::DescriptorDistance (ORBmatcher.cpp:1838-1854) / FORB::distance (DBoW2/FORB.cpp:82-102) on two 32-byte rows
__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

}  // namespace orbb
