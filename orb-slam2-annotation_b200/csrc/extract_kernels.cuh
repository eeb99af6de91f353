// extract_kernels.cuh — sm_100a kernels of the ORB extractor hot path.
//
//   level3_kernel    : (level_kernel3.cuh) one launch per pyramid level, grid = (tiles_x, tiles_y, frames).  A CTA of 4 (or 8) warps stages a
//                      120 x 32 (120 x 64) pixel tile of level L plus a 4-pixel halo in shared memory ONCE, by one bulk tensor copy (TMA), and
//                      produces from it (a) the 7x7 sigma-2 blurred level (ORBextractor.cpp:1097-1098), (b) the FAST-9-16 arc score, cell-local
//                      3x3 NMS and the raw corner records (ORBextractor.cpp:797-838), (c) its part of level L+1 by bilinear resize
//                      (ORBextractor.cpp:1136).  Level L is therefore read from HBM once.
//   octree_kernel    : DistributeOctTree (ORBextractor.cpp:541-770), one CTA per (level, frame).
//   orient_desc_kernel: IC_Angle (ORBextractor.cpp:79-106) + computeOrbDescriptor (:110-149) + the coordinate scaling
//                      and concatenation of operator() (:1084-1116), one warp per keypoint.
//
// Exact-arithmetic notes: integer fixed-point resize/blur, threshold-independent FAST score
// m = max over 9-arcs, float ops with explicit _rn intrinsics (no FMA contraction), round-half-even conversions.
#pragma once
#include "common.cuh"
#include "level_kernel3.cuh"

namespace orbb {

__constant__ int8_t c_pattern[1024];  // rBRIEF pattern (ORBextractor.cpp:152-410), x0,y0,x1,y1 per test
__constant__ int c_umax[16];          // circular patch row ends (ORBextractor.cpp:456-471)
__constant__ float c_atan[4];         // fastAtan2 polynomial p1,p3,p5,p7 (already scaled to degrees)

// =====================================================================================================
// cv::cvtColor RGB/BGR/RGBA/BGRA -> GRAY as Tracking::GrabImage* applies it in front of the extractor (Tracking.cpp:202-227,
// SURVEY 8(f)-4).  OpenCV 4.x 15-bit fixed point: (R*9798 + G*19235 + B*3735 + 2^14) >> 15.  A thread converts four pixels:
// word loads when the source row allows, one 32-bit store.  HBM-bound: 3-4 bytes read, 1 written per pixel.
// =====================================================================================================
__device__ __forceinline__ unsigned cvt_gray_px(unsigned c0, unsigned c1, unsigned c2, int rgb)
{
    const unsigned r = rgb ? c0 : c2, b = rgb ? c2 : c0;
    return (r * 9798u + c1 * 19235u + b * 3735u + 16384u) >> 15;
}

__global__ void __launch_bounds__(128) cvt_gray_kernel(const uint8_t* __restrict__ src, long long spitch, long long sfstride, int w, int ch, int rgb,
                                                       uint8_t* __restrict__ dst, long long dpitch, long long dfstride)
{
    const int x = 4 * (blockIdx.x * blockDim.x + threadIdx.x), y = blockIdx.y, f = blockIdx.z;
    if (x >= w) return;
    const uint8_t* s = src + (long long)f * sfstride + (long long)y * spitch + (long long)x * ch;
    uint8_t* d = dst + (long long)f * dfstride + (long long)y * dpitch + x;
    const int n = min(4, w - x);
    unsigned g[4] = {0u, 0u, 0u, 0u};
    if (n == 4 && ch == 4 && ((uintptr_t)s & 15) == 0) {
        const uint4 v = *reinterpret_cast<const uint4*>(s);
        const unsigned p[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; i++) g[i] = cvt_gray_px(p[i] & 255u, (p[i] >> 8) & 255u, (p[i] >> 16) & 255u, rgb);
    } else if (n == 4 && ch == 3 && ((uintptr_t)s & 3) == 0) {
        const unsigned a = *reinterpret_cast<const unsigned*>(s), b = *reinterpret_cast<const unsigned*>(s + 4), c = *reinterpret_cast<const unsigned*>(s + 8);
        g[0] = cvt_gray_px(a & 255u, (a >> 8) & 255u, (a >> 16) & 255u, rgb);
        g[1] = cvt_gray_px(a >> 24, b & 255u, (b >> 8) & 255u, rgb);
        g[2] = cvt_gray_px((b >> 16) & 255u, b >> 24, c & 255u, rgb);
        g[3] = cvt_gray_px((c >> 8) & 255u, (c >> 16) & 255u, c >> 24, rgb);
    } else {
        for (int i = 0; i < n; i++) g[i] = cvt_gray_px(s[i * ch], s[i * ch + 1], s[i * ch + 2], rgb);
    }
    if (n == 4 && ((uintptr_t)d & 3) == 0) *reinterpret_cast<unsigned*>(d) = g[0] | (g[1] << 8) | (g[2] << 16) | (g[3] << 24);
    else for (int i = 0; i < n; i++) d[i] = (uint8_t)g[i];
}

// Frames whose rows are contiguous on both sides (source pitch = w*CH, destination pitch = w, 16-byte aligned frame bases): the frame
// is one flat array.  A thread converts 16 pixels: CH 128-bit loads, one 128-bit store; a CTA covers 4096 pixels.
template <int CH>
__global__ void __launch_bounds__(256) cvt_gray_flat_kernel(const uint8_t* __restrict__ src, long long sfstride, int npix, int rgb,
                                                            uint8_t* __restrict__ dst, long long dfstride)
{
    const int f = blockIdx.y;
    const uint8_t* s = src + (long long)f * sfstride;
    uint8_t* d = dst + (long long)f * dfstride;
    const int p0 = 16 * (blockIdx.x * blockDim.x + threadIdx.x);
    if (p0 >= npix) return;
    if (p0 + 16 <= npix) {
        unsigned in[4 * CH];
#pragma unroll
        for (int i = 0; i < CH; i++) {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(s + (long long)p0 * CH) + i);
            in[4 * i] = v.x; in[4 * i + 1] = v.y; in[4 * i + 2] = v.z; in[4 * i + 3] = v.w;
        }
        unsigned out[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            unsigned g[4];
            if (CH == 4) {
#pragma unroll
                for (int i = 0; i < 4; i++) { const unsigned v = in[4 * j + i]; g[i] = cvt_gray_px(v & 255u, (v >> 8) & 255u, (v >> 16) & 255u, rgb); }
            } else {
                const unsigned a = in[3 * j], b = in[3 * j + 1], c = in[3 * j + 2];
                g[0] = cvt_gray_px(a & 255u, (a >> 8) & 255u, (a >> 16) & 255u, rgb);
                g[1] = cvt_gray_px(a >> 24, b & 255u, (b >> 8) & 255u, rgb);
                g[2] = cvt_gray_px((b >> 16) & 255u, b >> 24, c & 255u, rgb);
                g[3] = cvt_gray_px((c >> 8) & 255u, (c >> 16) & 255u, c >> 24, rgb);
            }
            out[j] = g[0] | (g[1] << 8) | (g[2] << 16) | (g[3] << 24);
        }
        *reinterpret_cast<uint4*>(d + p0) = make_uint4(out[0], out[1], out[2], out[3]);
    } else {
        for (int p = p0; p < npix; p++) d[p] = (uint8_t)cvt_gray_px(s[(long long)p * CH], s[(long long)p * CH + 1], s[(long long)p * CH + 2], rgb);
    }
}

// picks the flat kernel when the layout allows it, the row kernel otherwise
static inline void launch_cvt_gray(const uint8_t* src, long long spitch, long long sfstride, int w, int ht, int ch, int rgb,
                                   uint8_t* dst, long long dpitch, long long dfstride, int nf, cudaStream_t st)
{
    const long long npix = (long long)w * ht;
    const bool flat = spitch == (long long)w * ch && dpitch == w && npix < (1ll << 30) && (((uintptr_t)src | (uintptr_t)dst) & 15) == 0 &&
                      (nf == 1 || ((sfstride | dfstride) & 15) == 0);
    if (flat) {
        const dim3 grid((unsigned)((npix + 16 * 256 - 1) / (16 * 256)), nf);
        if (ch == 3) cvt_gray_flat_kernel<3><<<grid, 256, 0, st>>>(src, sfstride, (int)npix, rgb, dst, dfstride);
        else cvt_gray_flat_kernel<4><<<grid, 256, 0, st>>>(src, sfstride, (int)npix, rgb, dst, dfstride);
    } else {
        cvt_gray_kernel<<<dim3(ceil_div(w, 4 * 128), ht, nf), 128, 0, st>>>(src, spitch, sfstride, w, ch, rgb, dst, dpitch, dfstride);
    }
}

// =====================================================================================================
// cv::remap(src, dst, map1, map2, INTER_LINEAR), BORDER_CONSTANT 0 — stereo rectification in front of the extractor
// (Examples/Stereo/stereo_euroc.cc:174-185).  The map arrives in cv::remap's own fixed point (1/32 pixel); the
// bilinear weights (32-fx)(32-fy)*32 ... sum to 2^15; out = (sum w*p + 2^14) >> 15.  A thread produces four pixels and one
// 32-bit store; the map (8 B per pixel, shared by all frames of the batch) stays in L2.
// =====================================================================================================
// FLAT (dw % 4 == 0): the destination frame is walked as one array of 4-pixel groups (grid = groups x frames), so that CTAs are
// full whatever the row length; otherwise one CTA row per image row.
template <bool FLAT>
__global__ void __launch_bounds__(256) remap_kernel(const uint8_t* __restrict__ src, long long spitch, long long sfstride, int sw, int sh,
                                                    const int2* __restrict__ map, int dw, int dh, uint8_t* __restrict__ dst, long long dpitch, long long dfstride,
                                                    unsigned gpr_magic)
{
    int x, y, f;
    if (FLAT) {                                   // host guarantees: groups * (dw / 4) < 2^32 (magic division exact), frames < 2 GB
        const int g = blockIdx.x * blockDim.x + threadIdx.x, gpr = dw >> 2;
        y = (int)__umulhi((unsigned)g, gpr_magic); x = 4 * (g - y * gpr); f = blockIdx.y;
        if (y >= dh) return;
    } else {
        x = 4 * (blockIdx.x * blockDim.x + threadIdx.x); y = blockIdx.y; f = blockIdx.z;
        if (x >= dw) return;
    }
    const uint8_t* s = src + (long long)f * sfstride;
    uint8_t* d = FLAT ? dst + (long long)f * dfstride + (unsigned)(y * (int)dpitch + x) : dst + (long long)f * dfstride + (long long)y * dpitch + x;
    const int n = min(4, dw - x);
    unsigned g[4] = {0u, 0u, 0u, 0u};
    int2 m[4];
    const int2* mp = FLAT ? map + (unsigned)(y * dw + x) : map + (long long)y * dw + x;
    if (FLAT) {                                  // x % 4 == 0 and dw % 4 == 0: the four map entries are one aligned 32-byte run
        const int4 a = __ldg(reinterpret_cast<const int4*>(mp)), b = __ldg(reinterpret_cast<const int4*>(mp) + 1);
        m[0] = make_int2(a.x, a.y); m[1] = make_int2(a.z, a.w); m[2] = make_int2(b.x, b.y); m[3] = make_int2(b.z, b.w);
    } else {
#pragma unroll
        for (int i = 0; i < 4; i++) m[i] = i < n ? __ldg(mp + i) : make_int2(0, 0);
    }
    const unsigned sp32 = (unsigned)spitch;
    const bool small = FLAT;                     // 32-bit source offsets (checked by launch_remap)
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if (i >= n) break;
        const int ix = m[i].x >> 5, iy = m[i].y >> 5;
        const int fx = m[i].x & 31, fy = m[i].y & 31;
        if (small && (unsigned)ix < (unsigned)(sw - 1) && (unsigned)iy < (unsigned)(sh - 1)) {
            // all four taps inside the source: sum of (32-fx)(32-fy)*32 ... products, factored (exact integer arithmetic)
            const uint8_t* r0 = s + ((unsigned)iy * sp32 + (unsigned)ix);
            const int p00 = r0[0], p01 = r0[1], p10 = r0[sp32], p11 = r0[sp32 + 1];
            const int top = p00 * (32 - fx) + p01 * fx, bot = p10 * (32 - fx) + p11 * fx;
            g[i] = (unsigned)((top * (32 - fy) + bot * fy + 512) >> 10);
        } else {
            const int cx = max(-32768, min(32767, ix)), cy = max(-32768, min(32767, iy));
            const bool x0 = cx >= 0 && cx < sw, x1 = cx + 1 >= 0 && cx + 1 < sw, y0 = cy >= 0 && cy < sh, y1 = cy + 1 >= 0 && cy + 1 < sh;
            const uint8_t* r0 = s + (long long)cy * spitch + cx;
            const uint8_t* r1 = r0 + spitch;
            const int p00 = (x0 && y0) ? r0[0] : 0, p01 = (x1 && y0) ? r0[1] : 0, p10 = (x0 && y1) ? r1[0] : 0, p11 = (x1 && y1) ? r1[1] : 0;
            const int acc = ((32 - fx) * (32 - fy) * p00 + fx * (32 - fy) * p01 + (32 - fx) * fy * p10 + fx * fy * p11) * 32;
            g[i] = (unsigned)((acc + (1 << 14)) >> 15);
        }
    }
    if (n == 4 && ((uintptr_t)d & 3) == 0) *reinterpret_cast<unsigned*>(d) = g[0] | (g[1] << 8) | (g[2] << 16) | (g[3] << 24);
    else for (int i = 0; i < n; i++) d[i] = (uint8_t)g[i];
}

// Lane-consecutive variant (frames below 2 GB): a warp covers 128 consecutive destination pixels of the flattened frame, lane l
// takes pixels l, l+32, l+64, l+96, so that a warp's map loads are contiguous 256-byte runs and its source taps fall into one or
// two cache lines per load instruction (the L1 data pipe, not HBM, bounds this gather).  magic = ceil(2^32 / dw).
__global__ void __launch_bounds__(256) remap_lanes_kernel(const uint8_t* __restrict__ src, unsigned spitch, long long sfstride, int sw, int sh,
                                                          const int2* __restrict__ map, int dw, int npix, uint8_t* __restrict__ dst, unsigned dpitch,
                                                          long long dfstride, unsigned magic)
{
    const int lane = threadIdx.x & 31;
    const int base = 128 * (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) + lane;
    const uint8_t* s = src + (long long)blockIdx.y * sfstride;
    uint8_t* d = dst + (long long)blockIdx.y * dfstride;
    int2 m[4];
#pragma unroll
    for (int i = 0; i < 4; i++) m[i] = base + 32 * i < npix ? __ldg(map + base + 32 * i) : make_int2(-64, -64);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int p = base + 32 * i;
        if (p >= npix) break;
        const int ix = m[i].x >> 5, iy = m[i].y >> 5, fx = m[i].x & 31, fy = m[i].y & 31;
        unsigned g;
        if ((unsigned)ix < (unsigned)(sw - 1) && (unsigned)iy < (unsigned)(sh - 1)) {
            const uint8_t* r0 = s + ((unsigned)iy * spitch + (unsigned)ix);
            const int p00 = r0[0], p01 = r0[1], p10 = r0[spitch], p11 = r0[spitch + 1];
            const int top = p00 * (32 - fx) + p01 * fx, bot = p10 * (32 - fx) + p11 * fx;
            g = (unsigned)((top * (32 - fy) + bot * fy + 512) >> 10);
        } else {
            const int cx = max(-32768, min(32767, ix)), cy = max(-32768, min(32767, iy));
            const bool x0 = cx >= 0 && cx < sw, x1 = cx + 1 >= 0 && cx + 1 < sw, y0 = cy >= 0 && cy < sh, y1 = cy + 1 >= 0 && cy + 1 < sh;
            const uint8_t* r0 = s + (long long)cy * spitch + cx;
            const uint8_t* r1 = r0 + spitch;
            const int p00 = (x0 && y0) ? r0[0] : 0, p01 = (x1 && y0) ? r0[1] : 0, p10 = (x0 && y1) ? r1[0] : 0, p11 = (x1 && y1) ? r1[1] : 0;
            const int acc = ((32 - fx) * (32 - fy) * p00 + fx * (32 - fy) * p01 + (32 - fx) * fy * p10 + fx * fy * p11) * 32;
            g = (unsigned)((acc + (1 << 14)) >> 15);
        }
        const unsigned y = __umulhi((unsigned)p, magic), x = (unsigned)p - y * (unsigned)dw;
        d[y * dpitch + x] = (uint8_t)g;
    }
}

static inline void launch_remap(const uint8_t* src, long long spitch, long long sfstride, int sw, int sh, const int2* map, int dw, int dh,
                                uint8_t* dst, long long dpitch, long long dfstride, int nf, cudaStream_t st)
{
    const long long npix = (long long)dw * dh;
    if ((npix + 1024) * dw < (1ll << 32) && (long long)sh * spitch < (1ll << 31) && (long long)dh * dpitch < (1ll << 31)) {
        const unsigned magic = (unsigned)(((1ull << 32) + (unsigned)dw - 1) / (unsigned)dw);
        remap_lanes_kernel<<<dim3((unsigned)((npix + 1023) / 1024), nf), 256, 0, st>>>(src, (unsigned)spitch, sfstride, sw, sh, map, dw, (int)npix, dst,
                                                                                   (unsigned)dpitch, dfstride, magic);
    } else if ((dw & 3) == 0 && ((long long)(dw >> 2) * dh + 256) * (dw >> 2) < (1ll << 32) && (long long)sh * spitch < (1ll << 31) &&
               (long long)dh * dpitch < (1ll << 31)) {
        const long long groups = (long long)(dw >> 2) * dh;
        const unsigned magic = (unsigned)(((1ull << 32) + (unsigned)(dw >> 2) - 1) / (unsigned)(dw >> 2));
        remap_kernel<true><<<dim3((unsigned)((groups + 255) / 256), nf), 256, 0, st>>>(src, spitch, sfstride, sw, sh, map, dw, dh, dst, dpitch, dfstride, magic);
    } else
        remap_kernel<false><<<dim3(ceil_div(dw, 4 * 128), dh, nf), 128, 0, st>>>(src, spitch, sfstride, sw, sh, map, dw, dh, dst, dpitch, dfstride, 0u);
}

// =====================================================================================================
// Octree distribution (ORBextractor.cpp:541-770).  Order-free formulation: the reference's result depends on the
// keypoint ORDER only through "first maximum wins" inside a leaf (:751-767); since vKeys order is always the
// cell-row-major / pixel-row-major input order (stable partitions, :512-528), the winner is the max-response point
// with the smallest (cellRow, cellCol, y, x) key.  Node-list order (push_front of n1..n4, erase of the parent) and
// the (count, creation-sequence) sort of :690 are reproduced exactly with scans.
// =====================================================================================================
constexpr int NTO = 256;
constexpr int BIG_NODE = 512;

struct OctArgs {
    const LevelGeom* geom; int nlevels; int iniTh; int level0;     // the CTA's level = level0 + blockIdx.x
    const uint32_t* rec; long long rec_fstride; const int* rec_count;
    uint32_t* pts; long long pts_fstride;          // per level: 2 * rec_cap  (points + partition temp)
    uint4* nodes; long long nodes_fstride;         // per level: node_cap       {x0|y0<<16, x1|y1<<16, begin, count}
    int* iscr; long long iscr_fstride;             // per level: ISCR_PER_NODE * node_cap ints of scratch
    uint8_t* cellflag; long long cellflag_fstride; // per level: nCols*nRows bytes at cell_off
    const long long* node_off; const long long* cell_off;   // per-level offsets (elements)
    int4* out; long long out_fstride;              // per level: out_cap  {x, y, response, 0} in level coordinates
    int* out_count;                                // [frame][level]
    int* status;                                   // [frame] sticky error bits (1: raw overflow, 2: node pool, 4: out cap)
};
constexpr int ISCR_PER_NODE = 12;

__device__ __forceinline__ int warp_incl_scan(int v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
    return v;
}
// Exclusive scan over the block; returns the exclusive prefix and writes the block total.  All threads must call.
__device__ __forceinline__ int block_excl_scan(int v, int* total, int* s_warp /*[NTO/32 + 1]*/)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int inc = warp_incl_scan(v, lane);
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = lane < NTO / 32 ? s_warp[lane] : 0;
        int wi = warp_incl_scan(w, lane);
        if (lane < NTO / 32) s_warp[lane] = wi - w;
        if (lane == NTO / 32 - 1) s_warp[NTO / 32] = wi;
    }
    __syncthreads();
    const int r = s_warp[warp] + inc - v;
    *total = s_warp[NTO / 32];
    __syncthreads();
    return r;
}

struct Quad { int c[4]; };

__device__ __forceinline__ int pt_quadrant(uint32_t p, int midx, int midy)
{
    const int px = (p >> 8) & 0xfff, py = p >> 20;
    return (px < midx ? 0 : 1) + (py < midy ? 0 : 2);
}

// One warp divides one node in place (via tmp): returns the 4 child counts; children occupy consecutive
// sub-ranges [begin, begin+c0), ... in n1,n2,n3,n4 order.
__device__ Quad warp_divide(uint32_t* pts, uint32_t* tmp, uint4 nd, int lane, bool scatter)
{
    const int x0 = nd.x & 0xffff, y0 = nd.x >> 16, x1 = nd.y & 0xffff, y1 = nd.y >> 16;
    const int midx = x0 + ((x1 - x0 + 1) >> 1), midy = y0 + ((y1 - y0 + 1) >> 1);
    const int begin = nd.z, count = nd.w;
    Quad q; q.c[0] = q.c[1] = q.c[2] = q.c[3] = 0;
    for (int i = lane; i < ((count + 31) & ~31); i += 32) {
        int k = -1;
        if (i < count) k = pt_quadrant(pts[begin + i], midx, midy);
#pragma unroll
        for (int j = 0; j < 4; j++) q.c[j] += __popc(__ballot_sync(0xffffffffu, k == j));
    }
    if (!scatter) return q;
    int base[4] = {begin, begin + q.c[0], begin + q.c[0] + q.c[1], begin + q.c[0] + q.c[1] + q.c[2]};
    for (int i = lane; i < ((count + 31) & ~31); i += 32) {
        int k = -1; uint32_t p = 0;
        if (i < count) { p = pts[begin + i]; k = pt_quadrant(p, midx, midy); }
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const unsigned b = __ballot_sync(0xffffffffu, k == j);
            if (k == j) tmp[base[j] + __popc(b & ((1u << lane) - 1))] = p;
            base[j] += __popc(b);
        }
    }
    __syncwarp();
    for (int i = lane; i < count; i += 32) pts[begin + i] = tmp[begin + i];
    __syncwarp();
    return q;
}

// The whole block divides one big node (unordered scatter: point order inside a node is irrelevant).
__device__ Quad block_divide(uint32_t* pts, uint32_t* tmp, uint4 nd, int* s_cnt /*[8]*/)
{
    const int x0 = nd.x & 0xffff, y0 = nd.x >> 16, x1 = nd.y & 0xffff, y1 = nd.y >> 16;
    const int midx = x0 + ((x1 - x0 + 1) >> 1), midy = y0 + ((y1 - y0 + 1) >> 1);
    const int begin = nd.z, count = nd.w, lane = threadIdx.x & 31;
    if (threadIdx.x < 8) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < ((count + NTO - 1) / NTO) * NTO; i += NTO) {
        int k = -1;
        if (i < count) k = pt_quadrant(pts[begin + i], midx, midy);
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const unsigned b = __ballot_sync(0xffffffffu, k == j);
            if (lane == 0 && b) atomicAdd(&s_cnt[j], __popc(b));
        }
    }
    __syncthreads();
    Quad q; for (int j = 0; j < 4; j++) q.c[j] = s_cnt[j];
    __syncthreads();
    if (threadIdx.x == 0) { s_cnt[4] = begin; s_cnt[5] = begin + q.c[0]; s_cnt[6] = begin + q.c[0] + q.c[1]; s_cnt[7] = begin + q.c[0] + q.c[1] + q.c[2]; }
    __syncthreads();
    for (int i = threadIdx.x; i < ((count + NTO - 1) / NTO) * NTO; i += NTO) {
        int k = -1; uint32_t p = 0;
        if (i < count) { p = pts[begin + i]; k = pt_quadrant(p, midx, midy); }
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const unsigned b = __ballot_sync(0xffffffffu, k == j);
            int base = 0;
            if (lane == 0 && b) base = atomicAdd(&s_cnt[4 + j], __popc(b));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (k == j) tmp[base + __popc(b & ((1u << lane) - 1))] = p;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < count; i += NTO) pts[begin + i] = tmp[begin + i];
    __syncthreads();
    return q;
}

__device__ __forceinline__ uint4 make_child(uint4 p, int j, int begin, int count)
{
    const int x0 = p.x & 0xffff, y0 = p.x >> 16, x1 = p.y & 0xffff, y1 = p.y >> 16;
    const int midx = x0 + ((x1 - x0 + 1) >> 1), midy = y0 + ((y1 - y0 + 1) >> 1);
    const int cx0 = (j & 1) ? midx : x0, cx1 = (j & 1) ? x1 : midx;
    const int cy0 = (j & 2) ? midy : y0, cy1 = (j & 2) ? y1 : midy;
    return make_uint4((unsigned)cx0 | ((unsigned)cy0 << 16), (unsigned)cx1 | ((unsigned)cy1 << 16), (unsigned)begin, (unsigned)count);
}

__global__ void __launch_bounds__(NTO) octree_kernel(const __grid_constant__ OctArgs a)
{
    __shared__ int s_warp[NTO / 32 + 1];
    __shared__ int s_cnt[8];
    __shared__ int s_root[256];
    __shared__ int s_n, s_misc[8];

    const int L = a.level0 + blockIdx.x, f = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const LevelGeom g = a.geom[L];
    int* out_count = a.out_count + (long long)f * a.nlevels + L;
    int nraw = a.rec_count[(long long)f * a.nlevels + L];
    if (nraw > g.rec_cap) { if (tid == 0) atomicOr(a.status + f, 1); nraw = g.rec_cap; }
    if (g.nCols < 1 || nraw == 0) { if (tid == 0) *out_count = 0; return; }

    const uint32_t* __restrict__ rec = a.rec + (long long)f * a.rec_fstride + g.rec_off;
    uint32_t* pts = a.pts + (long long)f * a.pts_fstride + 2 * g.rec_off;
    uint32_t* tmp = pts + g.rec_cap;
    uint4* nodes = a.nodes + (long long)f * a.nodes_fstride + a.node_off[L];
    int* iscr = a.iscr + (long long)f * a.iscr_fstride + a.node_off[L] * ISCR_PER_NODE;
    const int NC = g.node_cap;
    int* listA = iscr;            // node ids front -> back
    int* listB = iscr + NC;
    int* expl = iscr + 2 * NC;    // expandable node ids of this pass, list order   /  phase 2: processing order
    int* nml = iscr + 3 * NC;     // kept (non-divided) node ids, list order
    int* cand = iscr + 4 * NC;    // phase-2 candidates (children with > 1 point)
    int* cq = iscr + 5 * NC;      // 4 child counts per divided node -> cq[4*e + j]   (4*NC ints)
    int* cbase = iscr + 9 * NC;   // exclusive prefix of non-empty children
    int* rank = iscr + 10 * NC;
    int* bigl = iscr + 11 * NC;   // big nodes of this pass
    uint8_t* cellflag = a.cellflag + (long long)f * a.cellflag_fstride + a.cell_off[L];
    int4* out = a.out + (long long)f * a.out_fstride + g.out_off;

    // ---- cell threshold fallback (ORBextractor.cpp:818-825): a cell uses the iniTh set if it is non-empty ----
    const int ncell = g.nCols * g.nRows;
    for (int i = tid; i < ncell; i += NTO) cellflag[i] = 0;
    if (tid == 0) s_n = 0;
    __syncthreads();
    for (int i = tid; i < nraw; i += NTO) {
        const uint32_t r = rec[i];
        const int m = r & 0xff, x = (r >> 8) & 0xfff, y = r >> 20;
        if (m > a.iniTh) cellflag[(int)__umulhi((unsigned)(y - 19), g.magicH) * g.nCols + (int)__umulhi((unsigned)(x - 19), g.magicW)] = 1;
    }
    __syncthreads();
    for (int i = tid; i < ((nraw + NTO - 1) / NTO) * NTO; i += NTO) {
        bool keep = false; uint32_t r = 0;
        if (i < nraw) {
            r = rec[i];
            const int m = r & 0xff, x = (r >> 8) & 0xfff, y = r >> 20;
            keep = (m > a.iniTh) || !cellflag[(int)__umulhi((unsigned)(y - 19), g.magicH) * g.nCols + (int)__umulhi((unsigned)(x - 19), g.magicW)];
        }
        const unsigned b = __ballot_sync(0xffffffffu, keep);
        int base = 0;
        if (lane == 0 && b) base = atomicAdd(&s_n, __popc(b));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (keep) {
            const int m = r & 0xff, x = (r >> 8) & 0xfff, y = r >> 20;
            tmp[base + __popc(b & ((1u << lane) - 1))] = ((unsigned)(y - 16) << 20) | ((unsigned)(x - 16) << 8) | (unsigned)m;
        }
    }
    __syncthreads();
    const int n = s_n;
    if (n == 0) { if (tid == 0) *out_count = 0; return; }

    // ---- roots (ORBextractor.cpp:545-587) ----
    const int nIni = g.nIni;
    for (int i = tid; i < nIni; i += NTO) s_root[i] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += NTO) {
        const int rx = (tmp[i] >> 8) & 0xfff;
        int ri = (int)__fdiv_rn((float)rx, g.hX);
        ri = min(ri, nIni - 1);
        atomicAdd(&s_root[ri], 1);
    }
    __syncthreads();
    if (tid == 0) {
        int acc = 0, size = 0;
        for (int i = 0; i < nIni; i++) {
            const int c = s_root[i];
            if (c > 0) {
                const int rx0 = (int)__fmul_rn(g.hX, (float)i), rx1 = (int)__fmul_rn(g.hX, (float)(i + 1));
                nodes[size] = make_uint4((unsigned)rx0, (unsigned)rx1 | ((unsigned)(g.h - 32) << 16), (unsigned)acc, (unsigned)c);
                listA[size] = size;
                size++;
            }
            s_root[i] = acc;          // becomes the scatter cursor
            acc += c;
        }
        s_misc[0] = size;
    }
    __syncthreads();
    for (int i = tid; i < n; i += NTO) {
        const uint32_t p = tmp[i];
        const int rx = (p >> 8) & 0xfff;
        int ri = (int)__fdiv_rn((float)rx, g.hX);
        ri = min(ri, nIni - 1);
        pts[atomicAdd(&s_root[ri], 1)] = p;
    }
    __syncthreads();

    int size = s_misc[0];
    int nodeCount = size;
    int* list = listA; int* nlist = listB;
    const int N = g.quota;
    int K = 0;                  // phase-2 candidate count
    bool finish = false, phase2 = false, poolFull = false;

    // ---- phase 1: divide every node with more than one point (ORBextractor.cpp:595-677) ----
    while (!finish && !phase2) {
        const int prevSize = size;
        // classify list entries (ordered compaction into expandable / kept)
        int E = 0, M = 0;
        for (int base = 0; base < size; base += NTO) {
            const int i = base + tid;
            int id = -1, flag = 0;
            if (i < size) { id = list[i]; flag = nodes[id].w > 1; }
            int tot;
            const int ex = block_excl_scan(flag, &tot, s_warp);
            if (i < size) { if (flag) expl[E + ex] = id; else nml[M + (i - base) - ex] = id; }
            const int chunk = min(NTO, size - base);
            E += tot; M += chunk - tot;
        }
        if (nodeCount + 4 * E > NC) { poolFull = true; break; }
        __syncthreads();
        // divide: big nodes by the whole block, the rest one warp each
        if (tid == 0) s_misc[1] = 0;
        __syncthreads();
        for (int e = tid; e < E; e += NTO)
            if (nodes[expl[e]].w >= BIG_NODE) bigl[atomicAdd(&s_misc[1], 1)] = e;
        __syncthreads();
        const int nbig = s_misc[1];
        for (int b = 0; b < nbig; b++) {
            const int e = bigl[b];
            const Quad q = block_divide(pts, tmp, nodes[expl[e]], s_cnt);
            if (tid == 0) { cq[4 * e] = q.c[0]; cq[4 * e + 1] = q.c[1]; cq[4 * e + 2] = q.c[2]; cq[4 * e + 3] = q.c[3]; }
        }
        for (int e = warp; e < E; e += NTO / 32) {
            const uint4 nd = nodes[expl[e]];
            if (nd.w >= BIG_NODE) continue;
            const Quad q = warp_divide(pts, tmp, nd, lane, true);
            if (lane == 0) { cq[4 * e] = q.c[0]; cq[4 * e + 1] = q.c[1]; cq[4 * e + 2] = q.c[2]; cq[4 * e + 3] = q.c[3]; }
        }
        __syncthreads();
        // children ids in creation order (parents front->back, n1..n4)
        int C = 0;
        for (int base = 0; base < E; base += NTO) {
            const int e = base + tid;
            int c = 0;
            if (e < E) c = (cq[4 * e] > 0) + (cq[4 * e + 1] > 0) + (cq[4 * e + 2] > 0) + (cq[4 * e + 3] > 0);
            int tot;
            const int ex = block_excl_scan(c, &tot, s_warp);
            if (e < E) cbase[e] = C + ex;
            C += tot;
        }
        if (tid == 0) { s_misc[2] = 0; s_misc[3] = 0; }
        __syncthreads();
        for (int e = tid; e < E; e += NTO) {
            const uint4 p = nodes[expl[e]];
            int k = cbase[e], begin = p.z, nexp = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int c = cq[4 * e + j];
                if (c > 0) {
                    const int id = nodeCount + k;
                    nodes[id] = make_child(p, j, begin, c);
                    nlist[C - 1 - k] = id;                 // push_front: newest child is the list front
                    if (c > 1) { nexp++; cand[atomicAdd(&s_misc[3], 1)] = id; }
                    k++; begin += c;
                }
            }
            if (nexp) atomicAdd(&s_misc[2], nexp);
        }
        for (int i = tid; i < M; i += NTO) nlist[C + i] = nml[i];
        __syncthreads();
        const int nToExpand = s_misc[2];
        K = s_misc[3];
        nodeCount += C;
        size = C + M;
        { int* t = list; list = nlist; nlist = t; }
        if (size >= N || size == prevSize) finish = true;
        else if (size + nToExpand * 3 > N) phase2 = true;
        __syncthreads();
    }

    // ---- phase 2: split the largest nodes first until N is reached (ORBextractor.cpp:678-745) ----
    while (!finish && !poolFull) {
        const int prevSize = size;
        if (nodeCount + 4 * K > NC) { poolFull = true; break; }
        // child counts of every candidate (no scatter yet)
        if (tid == 0) s_misc[1] = 0;
        __syncthreads();
        for (int k = tid; k < K; k += NTO)
            if (nodes[cand[k]].w >= BIG_NODE) bigl[atomicAdd(&s_misc[1], 1)] = k;
        __syncthreads();
        const int nbig = s_misc[1];
        for (int k = warp; k < K; k += NTO / 32) {
            const uint4 nd = nodes[cand[k]];
            const Quad q = warp_divide(pts, tmp, nd, lane, false);
            if (lane == 0) { cq[4 * k] = q.c[0]; cq[4 * k + 1] = q.c[1]; cq[4 * k + 2] = q.c[2]; cq[4 * k + 3] = q.c[3]; }
        }
        __syncthreads();
        // sort by (count, creation sequence) ascending; process from the back (ORBextractor.cpp:690-691)
        for (int k = tid; k < K; k += NTO) {
            const int idk = cand[k];
            const unsigned long long key = ((unsigned long long)nodes[idk].w << 32) | (unsigned)idk;
            int r = 0;
            for (int j = 0; j < K; j++) {
                const int idj = cand[j];
                r += (((unsigned long long)nodes[idj].w << 32) | (unsigned)idj) < key;
            }
            expl[K - 1 - r] = k;            // expl[t] = candidate processed t-th
        }
        __syncthreads();
        // how many get processed before lNodes.size() >= N (the `break` at :739-740)
        if (tid == 0) s_misc[4] = K;
        __syncthreads();
        int P = 0;
        for (int base = 0; base < K; base += NTO) {
            const int t = base + tid;
            int c = 0;
            if (t < K) { const int k = expl[t]; c = (cq[4 * k] > 0) + (cq[4 * k + 1] > 0) + (cq[4 * k + 2] > 0) + (cq[4 * k + 3] > 0); }
            int tot;
            const int ex = block_excl_scan(c, &tot, s_warp);
            if (t < K) {
                cbase[t] = P + ex;
                if (size + (P + ex + c) - (t + 1) >= N) atomicMin(&s_misc[4], t + 1);
            }
            P += tot;
        }
        __syncthreads();
        const int T = s_misc[4];           // nodes actually divided this round
        if (tid == 0) { s_misc[2] = 0; s_misc[3] = 0; }
        __syncthreads();
        // scatter the divided nodes
        for (int b = 0; b < nbig; b++) {
            const int k = bigl[b];
            // processed iff its position in expl is < T
            bool proc = false;
            for (int t = 0; t < T; t++) if (expl[t] == k) { proc = true; break; }
            if (proc) (void)block_divide(pts, tmp, nodes[cand[k]], s_cnt);
        }
        for (int t = warp; t < T; t += NTO / 32) {
            const uint4 nd = nodes[cand[expl[t]]];
            if (nd.w >= BIG_NODE) continue;
            (void)warp_divide(pts, tmp, nd, lane, true);
        }
        __syncthreads();
        int C = 0;
        if (T > 0) { const int k = expl[T - 1]; C = cbase[T - 1] + (cq[4 * k] > 0) + (cq[4 * k + 1] > 0) + (cq[4 * k + 2] > 0) + (cq[4 * k + 3] > 0); }
        // create children (creation order = processing order), mark parents erased
        for (int t = tid; t < T; t += NTO) {
            const int k = expl[t];
            const int pid = cand[k];
            const uint4 p = nodes[pid];
            int kk = cbase[t], begin = p.z;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int c = cq[4 * k + j];
                if (c > 0) {
                    const int id = nodeCount + kk;
                    nodes[id] = make_child(p, j, begin, c);
                    nlist[C - 1 - kk] = id;
                    if (c > 1) rank[atomicAdd(&s_misc[3], 1)] = id;     // next round's candidates
                    kk++; begin += c;
                }
            }
            nodes[pid].w = 0;                                           // erased (lNodes.erase(...->lit), :736)
        }
        __syncthreads();
        // surviving old nodes keep their relative order behind the new children
        int M = 0;
        for (int base = 0; base < size; base += NTO) {
            const int i = base + tid;
            int id = -1, flag = 0;
            if (i < size) { id = list[i]; flag = nodes[id].w > 0; }
            int tot;
            const int ex = block_excl_scan(flag, &tot, s_warp);
            if (flag) nlist[C + M + ex] = id;
            M += tot;
        }
        __syncthreads();
        const int newK = s_misc[3];
        for (int k = tid; k < newK; k += NTO) cand[k] = rank[k];
        nodeCount += C;
        size = C + M;
        K = newK;
        { int* t = list; list = nlist; nlist = t; }
        if (size >= N || size == prevSize) finish = true;
        __syncthreads();
    }
    if (poolFull && tid == 0) atomicOr(a.status + f, 2);

    // ---- best point per node, list order (ORBextractor.cpp:748-767) ----
    if (size > g.out_cap) { if (tid == 0) atomicOr(a.status + f, 4); size = g.out_cap; }
    for (int i = tid; i < size; i += NTO) {
        const uint4 nd = nodes[list[i]];
        int bm = -1; unsigned long long bkey = 0; uint32_t bp = 0;
        for (unsigned j = 0; j < nd.w; j++) {
            const uint32_t p = pts[nd.z + j];
            const int m = p & 0xff, x = ((p >> 8) & 0xfff) + 16, y = (p >> 20) + 16;
            const unsigned long long key = ((unsigned long long)((int)__umulhi((unsigned)(y - 19), g.magicH) * g.nCols + (int)__umulhi((unsigned)(x - 19), g.magicW)) << 24) | ((unsigned)y << 12) | (unsigned)x;
            if (m > bm || (m == bm && key < bkey)) { bm = m; bkey = key; bp = p; }
        }
        out[i] = make_int4((int)((bp >> 8) & 0xfff) + 16, (int)(bp >> 20) + 16, bm - 1, 0);
    }
    if (tid == 0) *out_count = size;
}

// =====================================================================================================
// Orientation + descriptor + final gather, one warp per keypoint.
// =====================================================================================================
struct DescArgs {
    const LevelGeom* geom; int nlevels;
    const uint8_t* src0; long long src0_pitch, src0_fstride;     // level 0 = the caller's image
    const uint8_t* pyr; long long pyr_fstride;                   // levels >= 1
    const uint8_t* blur; long long blur_fstride;
    const int4* lvl; long long lvl_fstride; const int* lvl_count;  // octree output
    orbx_keypoint* kps; uint8_t* desc; int* counts; int capacity;  // final output, frame f at f*capacity
    float factorPI;
    int kpw;                                                       // keypoints per warp: 32 for full batches, fewer when a launch would not fill the GPU
};

__device__ __forceinline__ float fast_atan2_deg(float y, float x)
{
    const float ax = fabsf(x), ay = fabsf(y);
    float a;
    if (ax >= ay) {
        const float c = __fdiv_rn(ay, __fadd_rn(ax, 2.2204460492503131e-16f));
        const float c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(c_atan[3], c2), c_atan[2]), c2), c_atan[1]), c2), c_atan[0]), c);
    } else {
        const float c = __fdiv_rn(ax, __fadd_rn(ay, 2.2204460492503131e-16f));
        const float c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(c_atan[3], c2), c_atan[2]), c2), c_atan[1]), c2), c_atan[0]), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// One warp owns 32 consecutive output slots.
//   Phase A (IC_Angle, ORBextractor.cpp:79-106): the warp visits its keypoints one by one.  The 31x31 window is read as
//            ALIGNED 32-bit words, 9 per row, three rows per warp load (lane = row%3 * 9 + word: coalesced, and the address
//            is a running pointer).  The circular mask is a packed byte compare (|u| <= umax[|v|]) and the moments come from
//            three IDP.4A per word: S = sum(p), T = sum(b*p) (b = byte 0..3), A = sum(it*p), from which
//            m10 = u0*S + T and m01 = 3*A + (row%3 - 15)*S per lane.  REDUX-reduced and parked in lane j.
//            The next keypoint's words are requested before the current ones are consumed.
//   Phase B: every lane finishes ITS keypoint (fastAtan2, cos/sin in double -> float: 32 keypoints per instruction).
//   Phase C (computeOrbDescriptor, :110-149): the 37x37 blurred window (10 aligned words per row) is staged in shared memory
//            with the same three-rows-per-load scheme, then lane = descriptor BYTE: 16 rotated pattern points per lane are
//            read from the staged window.  Again the next window is in flight while the current one is used.
constexpr int OD_WARPS = 8;
constexpr int OD_AW = 9, OD_AIT = 11;          // phase A: words per row, warp loads per window (33 >= 31 rows)
constexpr int OD_PW = 10, OD_PIT = 13;         // phase C: words per staged row (37 bytes + 3 alignment), warp loads (39 >= 37 rows)
constexpr int OD_PROWS = 37;

__device__ __forceinline__ unsigned od_dp4a_us(unsigned a, unsigned b, unsigned c)
{
    unsigned d; asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}

// NIT warp loads of three rows each: lane's word of row 3*it + r3.  Unconditional: rows past the window and the idle lanes'
// items stay inside the level (keypoints are >= 19 px from the border) or the buffer margin, and are masked by the consumer.
template <int NIT>
__device__ __forceinline__ void od_load_rows(unsigned (&w)[NIT], const uint8_t* q, unsigned p3)
{
#pragma unroll
    for (int it = 0; it < NIT; it++) w[it] = __ldg(reinterpret_cast<const unsigned*>(q + (unsigned long long)it * p3));
}

#ifndef OD_MINB
#define OD_MINB 3
#endif
__global__ void __launch_bounds__(32 * OD_WARPS, OD_MINB) orient_desc_kernel(const __grid_constant__ DescArgs a)
{
    __shared__ __align__(16) uint32_t patch_s[OD_WARPS][2][3 * OD_PIT * OD_PW + 2];      // two windows per warp: one in use, one in flight
    __shared__ uint32_t absu_s[40];            // |u0+b| packed per byte, indexed by u0 + 18
    const int f = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < 40) {
        const int u0 = (int)threadIdx.x - 18;
        absu_s[threadIdx.x] = (unsigned)abs(u0) | ((unsigned)abs(u0 + 1) << 8) | ((unsigned)abs(u0 + 2) << 16) | ((unsigned)abs(u0 + 3) << 24);
    }
    __syncthreads();
    const int wid = blockIdx.x * OD_WARPS + warp;
    const int base = wid * a.kpw;
    const int* cnt = a.lvl_count + (long long)f * a.nlevels;
    // own keypoint: (level, index) from the per-level counts
    const int slot = base + lane;
    int level = -1, idx = 0, acc = 0;
    for (int l = 0; l < a.nlevels; l++) {
        const int c = cnt[l];
        if (level < 0 && slot < acc + c) { level = l; idx = slot - acc; }
        acc += c;
    }
    const int total = acc;
    if (wid == 0 && lane == 0) a.counts[f] = total;
    const int nvalid = min(min(total, a.capacity) - base, a.kpw);   // keypoints handled by this warp
    if (nvalid <= 0) return;
    const bool mine = lane < nvalid;
    int kx = 0, ky = 0, kresp = 0, pitch = 0, bpitch = 0;
    const uint8_t* img = nullptr; const uint8_t* bimg = nullptr;
    float scale = 1.f, size = 0.f;
    if (mine) {
        const LevelGeom& g = a.geom[level];
        const int4 k = a.lvl[(long long)f * a.lvl_fstride + g.out_off + idx];
        kx = k.x; ky = k.y; kresp = k.z;
        if (level == 0) { img = a.src0 + (long long)f * a.src0_fstride; pitch = (int)a.src0_pitch; }
        else { img = a.pyr + (long long)f * a.pyr_fstride + g.off; pitch = g.pitch; }
        bimg = a.blur + (long long)f * a.blur_fstride + g.off; bpitch = g.pitch;      // level bases; the centre is (kx, ky)
        scale = g.scale; size = g.size;
    }
    // ---- phase A: intensity centroid ----
    int my_m10 = 0, my_m01 = 0;
    {
        const int r3 = lane / OD_AW, wc = lane - r3 * OD_AW;        // lanes 27..31 carry no item
        // umax of the lane's row in each of the 11 loads, as 0x80|umax bytes (four loads per register); 0x7f = nothing passes
        // (rows past the window, and lanes 27..31 whose rows belong to other lanes)
        unsigned umr[3] = {0u, 0u, 0u};
#pragma unroll
        for (int it = 0; it < OD_AIT; it++) {
            const int row = 3 * it + r3;
            const unsigned um = (r3 < 3 && row < 31) ? 0x80u | (unsigned)c_umax[min(abs(row - 15), 15)] : 0x7fu;
            umr[it >> 2] |= um << (8 * (it & 3));
        }
        unsigned wn[OD_AIT]; int cxn = 0, uan = 0; bool aln = false;
        const uint8_t* ibn = nullptr; int pn = 0, cyn = 0;
        // request keypoint j's window
        auto request = [&](int j) {
            ibn = (const uint8_t*)__shfl_sync(0xffffffffu, (unsigned long long)img, j);
            pn = __shfl_sync(0xffffffffu, pitch, j); cxn = __shfl_sync(0xffffffffu, kx, j); cyn = __shfl_sync(0xffffffffu, ky, j);
            aln = (((unsigned long long)ibn | (unsigned)pn) & 3) == 0;   // warp-uniform
            if (aln) {
                const int xs = (cxn - 15) & ~3;
                uan = xs - cxn;                                          // u of byte 0 of word 0 (-18 .. -15)
                od_load_rows<OD_AIT>(wn, ibn + (long long)(cyn - 15 + r3) * pn + xs + 4 * wc, 3u * (unsigned)pn);
            }
        };
        request(0);
        for (int j = 0; j < nvalid; j++) {
            unsigned w[OD_AIT];
#pragma unroll
            for (int it = 0; it < OD_AIT; it++) w[it] = wn[it];
            const bool al = aln; const int ua = uan, cx = cxn, cy = cyn, p = pn; const uint8_t* ib = ibn;
            if (j + 1 < nvalid) request(j + 1);
            int m10, m01;
            if (al) {
                const int u0 = ua + 4 * wc;
                const unsigned au = absu_s[u0 + 18];
                unsigned S = 0, T = 0, A = 0;
#pragma unroll
                for (int it = 0; it < OD_AIT; it++) {
                    const unsigned umb = __byte_perm(umr[it >> 2], 0u, 0x1111u * (it & 3));     // 0x80|umax in every byte
                    const unsigned ge = ((umb - au) >> 7) & 0x01010101u;                          // umax >= |u| per byte
                    const unsigned wm = w[it] & (ge * 0xffu);
                    S = od_dp4a_us(wm, 0x01010101u, S);
                    T = od_dp4a_us(wm, 0x03020100u, T);
                    A = od_dp4a_us(wm, 0x01010101u * (unsigned)it, A);
                }
                m10 = u0 * (int)S + (int)T;
                m01 = 3 * (int)A + (r3 - 15) * (int)S;
            } else {                                                    // unaligned caller image (level 0 only): bytes, lane = column
                const int u = lane - 15;
                const int dmax = lane < 31 ? c_umax[abs(u)] : -1;       // the disc is symmetric: |v| <= umax[|u|]
                const uint8_t* c = ib + (long long)cy * p + cx;
                int s = 0, sv = 0;
                for (int v = -15; v <= 15; v++)
                    if (abs(v) <= dmax) { const int val = c[v * p + u]; s += val; sv += v * val; }
                m10 = u * s; m01 = sv;
            }
            m10 = __reduce_add_sync(0xffffffffu, m10);
            m01 = __reduce_add_sync(0xffffffffu, m01);
            if (lane == j) { my_m10 = m10; my_m01 = m01; }
        }
    }
    // ---- phase B: per-lane angle and rotation ----
    const float angle = fast_atan2_deg((float)my_m01, (float)my_m10);
    const float ar = __fmul_rn(angle, a.factorPI);
    const float ca = (float)cos((double)ar), sb = (float)sin((double)ar);
    if (mine) {
        orbx_keypoint kp;
        kp.x = level ? __fmul_rn((float)kx, scale) : (float)kx;
        kp.y = level ? __fmul_rn((float)ky, scale) : (float)ky;
        kp.size = size; kp.angle = angle; kp.response = (float)kresp; kp.octave = level; kp.class_id = -1;
        a.kps[(long long)f * a.capacity + slot] = kp;
    }
    // ---- phase C: steered BRIEF; lane = descriptor byte ----
    float px[16], py[16];
#pragma unroll
    for (int i = 0; i < 16; i++) { px[i] = (float)c_pattern[lane * 32 + 2 * i]; py[i] = (float)c_pattern[lane * 32 + 2 * i + 1]; }
    uint8_t* dout = a.desc + ((long long)f * a.capacity + base) * 32 + lane;
    const int r3 = lane / OD_PW, wc = lane - r3 * OD_PW;           // lanes 30, 31 carry no item
    const bool lane_ok = lane < 3 * OD_PW;
    int con = 0; float cjn = 0.f, sjn = 0.f;
    // request keypoint j's blurred window into buffer `buf`: rows cy-18 .. cy+20, aligned words from (cx-18) & ~3 (the blurred
    // levels are pitch-aligned buffers of this library), global -> shared directly (LDGSTS), one commit group per window
    auto request = [&](int j, int buf) {
        const uint8_t* bb = (const uint8_t*)__shfl_sync(0xffffffffu, (unsigned long long)bimg, j);
        const int p = __shfl_sync(0xffffffffu, bpitch, j), cx = __shfl_sync(0xffffffffu, kx, j), cy = __shfl_sync(0xffffffffu, ky, j);
        cjn = __shfl_sync(0xffffffffu, ca, j); sjn = __shfl_sync(0xffffffffu, sb, j);
        const int xs = (cx - 18) & ~3;
        con = 18 * (4 * OD_PW) + (cx - xs);                         // window centre inside the staged rows
        if (lane_ok) {
            const uint8_t* q = bb + (long long)(cy - 18 + r3) * p + xs + 4 * wc;
            const unsigned p3 = 3u * (unsigned)p;
            const unsigned sa = (unsigned)__cvta_generic_to_shared(&patch_s[warp][buf][lane]);
#pragma unroll
            for (int it = 0; it < OD_PIT; it++)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa + 4u * (unsigned)(it * 3 * OD_PW)), "l"(q + (unsigned long long)it * p3) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    request(0, 0);
    for (int j = 0; j < nvalid; j++) {
        const uint8_t* c = reinterpret_cast<const uint8_t*>(patch_s[warp][j & 1]) + con;
        const float cj = cjn, sj = sjn;
        if (j + 1 < nvalid) { request(j + 1, (j + 1) & 1); asm volatile("cp.async.wait_group 1;" ::: "memory"); }
        else asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();                                                   // every lane's part of window j has landed
        int val = 0;
#pragma unroll
        for (int b = 0; b < 8; b++) {
            const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(px[2 * b], sj), __fmul_rn(py[2 * b], cj)));
            const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(px[2 * b], cj), __fmul_rn(py[2 * b], sj)));
            const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(px[2 * b + 1], sj), __fmul_rn(py[2 * b + 1], cj)));
            const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(px[2 * b + 1], cj), __fmul_rn(py[2 * b + 1], sj)));
            const int t0 = c[r0 * (4 * OD_PW) + c0], t1 = c[r1 * (4 * OD_PW) + c1];
            val |= (t0 < t1) << b;
        }
        dout[j * 32] = (uint8_t)val;
        __syncwarp();                                                   // window j is free for request j + 2
    }
}

}  // namespace orbb
