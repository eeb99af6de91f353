// level_kernel3.cuh — fused per-pyramid-level kernel, third design (sm_100a).
//
// One launch per level, grid = (tiles_x, tiles_y, frames).  A CTA of NW warps owns a tile of 120 x (8*NW) pixels of
// level L.  The tile plus a 4-pixel halo is staged ONCE in shared memory as 32 words x (8*NW + 8) rows, so that
//       lane  <->  32-bit word column (4 pixels),      warp  <->  group of 8 rows.
// From the staged tile the CTA produces
//   (a) the blurred level   cv::GaussianBlur 7x7 sigma 2 (ORBextractor.cpp:1097-1098).  The 2-D fixed-point convolution is
//       exact integer arithmetic (no intermediate rounding), so it is evaluated VERTICAL FIRST on bytes unpacked to u16x2
//       pairs (one IMAD = two pixels; sums stay below 2^16), then horizontally with IDP.2A on the Q8.8 pairs; the left /
//       right neighbour columns come from warp shuffles, never from shared memory.
//   (b) FAST-9-16 candidates  cv::FAST per 30-px cell + 3x3 NMS (ORBextractor.cpp:797-838): word-SIMD quick reject
//       (VABSDIFF4 on the 4 axis + 4 diagonal ring pixels), warp-compacted candidate queue, exact threshold-independent
//       arc score on packed u16x2 min/max (VIMNMX3), cell-local NMS, one global atomic per CTA.
//   (c) the tile's part of level L+1   cv::resize INTER_LINEAR (ORBextractor.cpp:1136): a lane owns one DESTINATION word;
//       horizontal interpolations are cached across destination rows; one 32-bit store per 4 pixels.
// Level L is read from HBM once; every store is a coalesced 32-bit word.  Bit-exact with the oracle.
#pragma once
#include "common.cuh"
#include <cuda.h>                             // CUtensorMap (the driver entry point is fetched at run time, no -lcuda)

namespace orbb {

constexpr int L3_LANES = 32;                 // OW = 30: staged word columns halo word | 30 owned words | halo word, one per lane
constexpr int L3_OWNW = 30;                  // owned words per tile row (default tile)
constexpr int L3_TWPX = 4 * L3_OWNW;         // 120 owned pixels per tile row
constexpr int L3_RG = 8;                     // rows per warp
// Tile width variants (template parameter OW = owned words per tile row):
//   OW = 30: 120-pixel tiles, lane l <-> staged word l (lanes 0 and 31 carry the halo words), shared pitch 36 words (the TMA box is
//            144 B wide because its first column must be 16-byte aligned in global memory; tile word c sits at cb + 1 + c, cb = 1 or 3).
//   OW = 32: 128-pixel tiles, lane l <-> owned word l, the two halo words (and the FAST ring pixels -1 and 128) are handled by small
//            extra passes; shared pitch 40 words (box 160 B).  Chosen by the host for the levels where it saves a tile column
//            (640 -> 5 instead of 6, 370 -> 3 instead of 4).  Measured (512 C1 frames): level 0 0.488 -> 0.467 ms, level 3 0.226 ->
//            0.219 ms - a full 128-pixel tile costs about 10 % more than a 120-pixel one, and the partly filled last column it
//            removes was cheap in the candidate-dependent phases, so the level sequence gains 1.5 %, not the 7 % the lane count suggests.
__host__ __device__ constexpr int l3_sp(int ow) { return ow == 32 ? 40 : 36; }           // shared tile pitch in words
__host__ __device__ constexpr int l3_scp(int ow) { return ow == 32 ? 136 : 128; }        // score map pitch in bytes (byte column = staged tile byte column)
__host__ __device__ constexpr int l3_scp_shift(int ow) { return ow == 32 ? 8 : 7; }      // queue entries are (score-map row << shift) | byte column

struct Level3Args {
    CUtensorMap tmap;        // level L as a 3-D tensor (x, y, frame), box 144 B x (tile rows + 8) x 1; valid when use_tma
    int use_tma;             // 0: source pitch / base not 16-byte aligned -> the tile is staged with ordinary loads
    int* status;             // [frame] sticky error bits (8: TMA transaction never completed)
    const uint8_t* src; long long src_pitch, src_fstride;   // level L (external input image when L == 0)
    int w, h;
    uint8_t* blur; long long blur_pitch, blur_fstride;      // blurred level L out
    uint8_t* dst; long long dst_pitch, dst_fstride;         // level L+1 out (null: last level, or resize done elsewhere)
    const int2* xt;          // per destination x (padded to a multiple of 4): {source x, a0 | a1 << 16}
    const int2* yt;          // per destination y: {source y, b0 | b1 << 16}
    const int* xws;          // [tiles_x + 1] first destination WORD owned by each tile column
    const int* yrs;          // [ceil(h / 8) + 1] first destination row whose source row lies in each 8-row group
    int fast_on, wCell, hCell, minTh;
    unsigned qmask, qadd;                                   // quick-reject constants: ~t' and 0x80 - 2^(k-1) per byte, t' = 2^k - 1 <= minTh
    unsigned magicW, magicH;                                // ceil(2^32 / wCell), ceil(2^32 / hCell)
    uint32_t* rec; int rec_cap; long long rec_fstride;      // raw FAST records  y<<20 | x<<8 | m
    int* rec_count; int rec_count_stride;
    int q[7];                                               // Q8 blur kernel
    unsigned kq[8];                                         // horizontal-pass tap words KA..KH (l3_blur_taps), packed by the host
};

// Tap words of the horizontal blur pass: byte pairs multiply the (even | odd) u16 halves of the left / own / right word.
inline void l3_blur_taps(const int* q, unsigned* k)
{
    auto pack = [](int b0, int b1, int b2, int b3) { return (unsigned)b0 | ((unsigned)b1 << 8) | ((unsigned)b2 << 16) | ((unsigned)b3 << 24); };
    const int q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
    k[0] = pack(q0, q2, 0, q1); k[1] = pack(q3, q1, q2, q0); k[2] = pack(0, q0, 0, q1); k[3] = pack(q2, q2, q3, q1);
    k[4] = pack(q0, 0, q1, 0);  k[5] = pack(q1, q3, q2, q2); k[6] = pack(q0, q2, q1, q3); k[7] = pack(q2, q0, q1, 0);
}

__device__ __forceinline__ int l3_reflect101(int p, int n)
{
    if (p < 0) p = -p;
    if (p >= n) p = 2 * (n - 1) - p;
    return min(max(p, 0), n - 1);
}

// Threshold-independent FAST arc score m = max over the 16 arcs of 9 contiguous ring pixels of
// max(min(v - p_k), min(p_k - v)); cv::FAST's response is m - 1 and a pixel is a corner at threshold t iff m > t.
// Both polarities ride in one register as biased u16 halves: lo = 256 + v - p_k, hi = 256 - v + p_k.
template <int P>
__device__ __forceinline__ int l3_arc_score(const uint8_t* c)
{
    const unsigned v = c[0];
    const unsigned av = (256u + v) + ((256u - v) << 16);
    unsigned X[16];
    X[0] = c[3 * P] * 65535u + av;       X[1] = c[3 * P + 1] * 65535u + av;   X[2] = c[2 * P + 2] * 65535u + av;
    X[3] = c[P + 3] * 65535u + av;       X[4] = c[3] * 65535u + av;           X[5] = c[-P + 3] * 65535u + av;
    X[6] = c[-2 * P + 2] * 65535u + av;  X[7] = c[-3 * P + 1] * 65535u + av;  X[8] = c[-3 * P] * 65535u + av;
    X[9] = c[-3 * P - 1] * 65535u + av;  X[10] = c[-2 * P - 2] * 65535u + av; X[11] = c[-P - 3] * 65535u + av;
    X[12] = c[-3] * 65535u + av;         X[13] = c[P - 3] * 65535u + av;      X[14] = c[2 * P - 2] * 65535u + av;
    X[15] = c[3 * P - 1] * 65535u + av;
    unsigned a3[16];
#pragma unroll
    for (int k = 0; k < 16; k++) a3[k] = __vimin3_u16x2(X[k], X[(k + 1) & 15], X[(k + 2) & 15]);
    // arcs k and k + 3 share a3[k+3], a3[k+6]:  max(a9[k], a9[k+3]) = min(a3[k+3], a3[k+6], max(a3[k], a3[k+9])); the eight pairs
    // (k, k + 3), k = 0, 6, 12, 2, 8, 14, 4, 10 cover the sixteen arcs
    unsigned pv[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int k = (6 * i) & 15;
        pv[i] = __vimin3_u16x2(a3[(k + 3) & 15], a3[(k + 6) & 15], __vmaxu2(a3[k], a3[(k + 9) & 15]));
    }
    unsigned best = __vimax3_u16x2(pv[0], pv[1], pv[2]);
    best = __vimax3_u16x2(best, pv[3], pv[4]);
    best = __vimax3_u16x2(best, pv[5], pv[6]);
    best = __vmaxu2(best, pv[7]);
    return (int)max(best & 0xffffu, best >> 16) - 256;
}

// LEGACY: the Q8 kernel of OpenCV <= 3.3 sums to 257, so the rounded result needs a clamp (default kernel: it is byte 2 of the sum).
// SPLIT (launches of at most a CTA or two per SM: one frame, a stereo pair): the tile keeps its NW row groups but the CTA has 2 * NW
// warps - warps 0 .. NW-1 run the FAST reject while warps NW .. 2NW-1 run blur + resize on the same staged tile, and all 2 * NW share
// the scoring and the NMS.  Same results; the tile's serial chain is about 40 % shorter, which is what a lone CTA per SM is bound by.
template <int NW, bool LEGACY, bool SPLIT = false, int OW = 30>
#ifndef L3_MINB
#define L3_MINB 8
#endif
__global__ void __launch_bounds__(SPLIT ? 64 * NW : 32 * NW, (NW == 8 || SPLIT) ? L3_MINB / 2 : L3_MINB) level3_kernel(const __grid_constant__ Level3Args a)
{
    constexpr int TH = L3_RG * NW, SMH = TH + 8, SCH = TH + 2, NWT = SPLIT ? 2 * NW : NW, NT = 32 * NWT;
    constexpr int L3_SP = l3_sp(OW), L3_SPB = 4 * L3_SP, L3_SCP = l3_scp(OW), SCSH = l3_scp_shift(OW), TWPX = 4 * OW;
    constexpr int L3_TWPX = TWPX, L3_OWNW = OW;          // (shadow the namespace constants of the default tile)
    constexpr int NSW = OW + 2;                          // staged words per row that are used
    constexpr int QCAP = SCH * (TWPX + 2);               // every scored pixel a candidate (uniform-noise images)
    constexpr unsigned FULL = 0xffffffffu;
    __shared__ __align__(128) uint32_t tws[(SMH + NW - 2) * L3_SP];   // tile + halo, 4 pixels per word (TMA destination: SMH rows); the
                                                                      // FAST row window of the last warp reads NW - 2 rows past it (masked)
    __shared__ __align__(8) unsigned long long mbar;      // completion barrier of the bulk tensor copy
    __shared__ __align__(16) uint8_t sc[SCH * L3_SCP];   // arc scores (> minTh) of the tile's pixels + 1-px ring
    __shared__ uint16_t queue[QCAP];
    __shared__ int s_qn;
    __shared__ uint2 hsum[OW == 32 ? NW * L3_RG * 2 : 1];  // OW = 32: vertical blur sums (even | odd pairs) of the two halo words, [row][side]

    const int tid = threadIdx.x, lane = tid & 31, wraw = tid >> 5;
    const int warp = SPLIT ? (wraw & (NW - 1)) : wraw;       // row group of this warp
    const bool role_fast = !SPLIT || wraw < NW, role_blur = !SPLIT || wraw >= NW;
    const int lw = OW == 32 ? lane + 1 : lane;               // index of this lane's word in a staged row (0 = the left halo word)
    const int f = blockIdx.z;
    const int x0 = blockIdx.x * L3_TWPX, y0 = blockIdx.y * TH;
    const uint8_t* __restrict__ src = a.src + (long long)f * a.src_fstride;
    // the box starts at the 16-byte boundary below x0 - 4: the tile's first word is word cb of each staged row
    const int cb = ((x0 - 4) & 15) >> 2;
    uint32_t* const tw = tws + cb;                          // tile word (r, c) = tw[r * L3_SP + c]
    const uint8_t* tb = reinterpret_cast<const uint8_t*>(tw);
    if (tid == 0) s_qn = 0;
    // SPLIT launches are chained by programmatic dependent launch: the next level's CTAs may become resident now and run their prologue;
    // they wait below, before the first read of this kernel's output (no-ops when the launch carries no such attribute)
    if (SPLIT) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

#pragma unroll
    for (int i = 0; i < (SCH * L3_SCP / 16 + NT - 1) / NT; i++)                                  // visible after the next barrier
        if (i * NT + tid < SCH * L3_SCP / 16) reinterpret_cast<uint4*>(sc)[i * NT + tid] = make_uint4(0u, 0u, 0u, 0u);
    if (SPLIT) asm volatile("griddepcontrol.wait;" ::: "memory");      // the previous level (this level's source image) is complete and visible

    // ---- 1. stage tile + halo (128 bytes x SMH rows, lane = word column).  Normal case: ONE bulk tensor copy (TMA) issued by
    //         thread 0 lands the box [(x0-4) & ~15, +144) x [y0-4, y0+TH+4) of frame f in shared memory (out-of-image bytes arrive as
    //         zeros) and signals an mbarrier; border tiles then mirror the few missing rows / columns in place (REFLECT_101, what
    //         the blur of a cloned level sees).  Fallback (source not 16-byte aligned): ordinary loads, all issued before the
    //         first store. ----
    if (a.use_tma) {
        const unsigned mb = (unsigned)__cvta_generic_to_shared(&mbar), dstp = (unsigned)__cvta_generic_to_shared(tws);
        if (tid == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(SMH * L3_SPB) : "memory");
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(dstp), "l"(reinterpret_cast<unsigned long long>(&a.tmap)), "r"((x0 - 4) & ~15), "r"(y0 - 4), "r"(f), "r"(mb) : "memory");
        }
        __syncthreads();                                          // the barrier is initialised before anyone polls it
        unsigned done = 0;
        for (int spin = 0; spin < (1 << 22) && !done; spin++)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(mb) : "memory");
        if (!done && tid == 0) atomicOr(a.status + f, 8);
        const bool ytop = y0 < 4, ybot = y0 + TH + 4 > a.h, xl = x0 < 4, xr = x0 + L3_TWPX + 4 > a.w;
        if (ytop || ybot || xl || xr) {                           // block-uniform: border tile
            if (ytop || ybot) {                                   // rows outside the image <- their mirror rows (which are inside)
                for (int r = wraw; r < SMH; r += NWT) {
                    const int gy = y0 - 4 + r;
                    if (gy < 0 || gy >= a.h) {
                        const int rs = min(max(l3_reflect101(gy, a.h) - (y0 - 4), 0), SMH - 1);
                        tw[r * L3_SP + lane] = tw[rs * L3_SP + lane];
                        if (OW == 32 && lane < 2) tw[r * L3_SP + 32 + lane] = tw[rs * L3_SP + 32 + lane];
                    }
                }
                __syncthreads();
            }
            if (xl || xr) {
                uint8_t* tbw = reinterpret_cast<uint8_t*>(tw);
                for (int r = tid; r < SMH; r += NT) {
                    if (xl) tw[r * L3_SP] = __byte_perm(tw[r * L3_SP + 1], tw[r * L3_SP + 2], 0x1234);      // pixels -4..-1 <- 4, 3, 2, 1
                    if (xr) {
#pragma unroll
                        for (int b = 0; b < 4; b++) {                                                     // pixels w..w+3 <- w-2 .. w-5
                            const int di = a.w + b - (x0 - 4), si = l3_reflect101(a.w + b, a.w) - (x0 - 4);
                            if (di >= 0 && di < 4 * NSW) tbw[r * L3_SPB + di] = tbw[r * L3_SPB + max(si, 0)];
                        }
                    }
                }
            }
            __syncthreads();
        }
    } else {
        constexpr int RPW = SMH / NW;                             // staged rows per warp (SMH is a multiple of NW)
        const bool word_ok = ((a.src_pitch & 3) == 0) && ((reinterpret_cast<unsigned long long>(src) & 3) == 0);
        const unsigned pitch = (unsigned)a.src_pitch;             // one frame is < 4 GB: 32-bit offsets
        // staged word column `col` of the rows warp, warp + NW, ...: every lane issues the same loads from a safe address and selects
        // afterwards, so the loads go out back to back
        auto stage_col = [&](int col) {
            const int gx = x0 - 4 + 4 * col;
            const bool inside = word_ok && gx >= 0 && gx + 3 < a.w;
            const bool bytes = !inside && gx < a.w + 4;           // edge lanes: 4 reflected byte loads per row
            const unsigned xw = inside ? (unsigned)gx : 0u;
            const unsigned xb0 = bytes ? l3_reflect101(gx, a.w) : 0, xb1 = bytes ? l3_reflect101(gx + 1, a.w) : 0;
            const unsigned xb2 = bytes ? l3_reflect101(gx + 2, a.w) : 0, xb3 = bytes ? l3_reflect101(gx + 3, a.w) : 0;
            unsigned v[RPW];
#pragma unroll
            for (int i = 0; i < RPW; i++) {
                const uint8_t* row = src + (unsigned)l3_reflect101(y0 - 4 + warp + NW * i, a.h) * pitch;
                unsigned vw = 0;
                if (word_ok) vw = *reinterpret_cast<const uint32_t*>(row + xw);
                const unsigned vb = (unsigned)row[xb0] | ((unsigned)row[xb1] << 8) | ((unsigned)row[xb2] << 16) | ((unsigned)row[xb3] << 24);
                v[i] = inside ? vw : (bytes ? vb : 0u);
            }
#pragma unroll
            for (int i = 0; i < RPW; i++) tw[(warp + NW * i) * L3_SP + col] = v[i];
        };
        if (!SPLIT || wraw < NW) {                                // (SPLIT: the first NW warps stage the tile)
            stage_col(lane);
            if (OW == 32 && lane < 2) stage_col(32 + lane);
        }
        __syncthreads();
    }

    const int ry0 = L3_RG * warp;                 // this warp's rows inside the tile
    const bool rows_live = y0 + ry0 < a.h;

    // ---- 2. blur: vertical pass on u16x2 pairs (even pixels | odd pixels), horizontal pass with IDP.2A ----
    if (rows_live && role_blur) {
        const unsigned q0 = a.q[0], q1 = a.q[1], q2 = a.q[2], q3 = a.q[3];      // symmetric kernel (checked by the host)
        if (OW == 32) {
            // the vertical sums of the two halo words, which no lane owns: lanes 0..15 take one (row, side) each of the warp's 8 rows
            if (lane < 2 * L3_RG) {
                const int j = lane >> 1, side = lane & 1;
                const uint32_t* hp = tw + (ry0 + 1 + j) * L3_SP + (side ? OW + 1 : 0);
                unsigned e[7], o[7];
#pragma unroll
                for (int i = 0; i < 7; i++) { const unsigned w = hp[i * L3_SP]; e[i] = w & 0x00ff00ffu; o[i] = __byte_perm(w, 0u, 0x4341); }
                hsum[(ry0 + j) * 2 + side] = make_uint2(q3 * e[3] + q2 * (e[2] + e[4]) + q1 * (e[1] + e[5]) + q0 * (e[0] + e[6]),
                                                        q3 * o[3] + q2 * (o[2] + o[4]) + q1 * (o[1] + o[5]) + q0 * (o[0] + o[6]));
            }
            __syncwarp();
        }
        unsigned E[14], O[14];
#pragma unroll
        for (int i = 0; i < 14; i++) {
            const unsigned w = tw[(ry0 + 1 + i) * L3_SP + lw];
            E[i] = w & 0x00ff00ffu;
            O[i] = __byte_perm(w, 0u, 0x4341);
        }
        const unsigned KA = a.kq[0], KB = a.kq[1], KC = a.kq[2], KD = a.kq[3], KE = a.kq[4], KF = a.kq[5], KG = a.kq[6], KH = a.kq[7];
        const uint2* const hedge = hsum + ry0 * 2 + (lane == 31 ? 1 : 0);      // (OW = 32) this lane's side of the halo sums, row j at [2 * j]
        const int gx = x0 - 4 + 4 * lw;
        const bool store_lane = (OW == 32 || (lane >= 1 && lane <= L3_OWNW)) && gx < a.w;
        uint8_t* __restrict__ bp = a.blur + (long long)f * a.blur_fstride + (long long)(y0 + ry0) * a.blur_pitch + gx;
        const int nrows = store_lane ? min(L3_RG, a.h - (y0 + ry0)) : 0;
#pragma unroll
        for (int j = 0; j < L3_RG; j++) {
            const unsigned VE = q3 * E[j + 3] + q2 * (E[j + 2] + E[j + 4]) + q1 * (E[j + 1] + E[j + 5]) + q0 * (E[j] + E[j + 6]);
            const unsigned VO = q3 * O[j + 3] + q2 * (O[j + 2] + O[j + 4]) + q1 * (O[j + 1] + O[j + 5]) + q0 * (O[j] + O[j + 6]);
            unsigned lE = __shfl_up_sync(FULL, VE, 1), lO = __shfl_up_sync(FULL, VO, 1);
            unsigned rE = __shfl_down_sync(FULL, VE, 1), rO = __shfl_down_sync(FULL, VO, 1);
            if (OW == 32) {                                          // the edge lanes' outer neighbours are the halo words: one broadcast load, four selects
                const uint2 t = hedge[2 * j];
                lE = lane == 0 ? t.x : lE; lO = lane == 0 ? t.y : lO;
                rE = lane == 31 ? t.x : rE; rO = lane == 31 ? t.y : rO;
            }
            // taps: q0..q6 with q4=q2, q5=q1, q6=q0; pairs are (even, odd) = (V0|V2<<16, V1|V3<<16) of the left / own / right word
            unsigned r0 = __dp2a_lo(lO, KA, 32768u); r0 = __dp2a_hi(lE, KA, r0); r0 = __dp2a_lo(VE, KB, r0); r0 = __dp2a_hi(VO, KB, r0);
            unsigned r1 = __dp2a_lo(lE, KC, 32768u); r1 = __dp2a_hi(lO, KC, r1); r1 = __dp2a_lo(VE, KD, r1); r1 = __dp2a_hi(VO, KD, r1);
            r1 = __dp2a_lo(rE, KE, r1);
            unsigned r2 = __dp2a_lo(lO, KC, 32768u); r2 = __dp2a_lo(VE, KF, r2); r2 = __dp2a_hi(VO, KF, r2); r2 = __dp2a_hi(rE, KE, r2);
            r2 = __dp2a_lo(rO, KE, r2);
            unsigned r3 = __dp2a_lo(VE, KG, 32768u); r3 = __dp2a_hi(VO, KG, r3); r3 = __dp2a_lo(rE, KH, r3); r3 = __dp2a_hi(rO, KH, r3);
            unsigned o;
            if (!LEGACY) {          // sums < 2^24: the rounded result is byte 2
                o = __byte_perm(__byte_perm(r0, r1, 0x0062), __byte_perm(r2, r3, 0x0062), 0x5410);
            } else {
                o = min(r0 >> 16, 255u) | (min(r1 >> 16, 255u) << 8) | (min(r2 >> 16, 255u) << 16) | (min(r3 >> 16, 255u) << 24);
            }
            if (j < nrows) *reinterpret_cast<uint32_t*>(bp) = o;
            bp += a.blur_pitch;
        }
    }

    // ---- 3. bilinear resize: lane = one destination word whose first source column lies in this tile; the warp walks the 9
    //         staged source rows of its group once (horizontal pass per source row, unrolled), emitting a destination row
    //         whenever its two source rows are at hand ----
    if (a.dst && rows_live && role_blur) {
        uint8_t* __restrict__ dst = a.dst + (long long)f * a.dst_fstride;
        const int k0 = a.xws[blockIdx.x], k1 = a.xws[blockIdx.x + 1];
        const int g8 = (y0 + ry0) >> 3;
        const int dy0 = a.yrs[g8], dy1 = a.yrs[g8 + 1];
        int2 eyl = make_int2(0x7fffffff, 0);                     // this group's yt entries, one per lane (at most 8 rows)
        if (dy0 + lane < dy1) eyl = a.yt[dy0 + lane];
        for (int kb = k0; kb < k1; kb += 32) {
            const bool active = kb + lane < k1;
            const int k = active ? kb + lane : k1 - 1;
            // the 4 pixels of a destination word read source bytes s0 .. s0+5 (scale <= 1.25): an 8-byte window of the staged row
            const int4 e01 = *reinterpret_cast<const int4*>(a.xt + 4 * k), e23 = *reinterpret_cast<const int4*>(a.xt + 4 * k + 2);
            const int off0 = e01.x - (x0 - 4);                   // byte offset of the first source pixel inside a staged row
            const unsigned sh = 8u * (unsigned)(off0 & 3);
            const unsigned r1 = (unsigned)(e01.z - e01.x), r2 = (unsigned)(e23.x - e01.x), r3 = (unsigned)(e23.z - e01.x);
            const unsigned sel0 = 0x10u, sel1 = ((r1 + 1) << 4) | r1, sel2 = ((r2 + 1) << 4) | r2, sel3 = ((r3 + 1) << 4) | r3;
            const unsigned c0 = (unsigned)e01.y, c1 = (unsigned)e01.w, c2 = (unsigned)e23.y, c3 = (unsigned)e23.w;   // a0 | a1 << 16
            const uint32_t* p0 = tw + (ry0 + 4) * L3_SP + (off0 >> 2);
            uint8_t* dp = dst + (long long)dy0 * a.dst_pitch + 4 * k;
            int dy = dy0;
            int sy = __shfl_sync(FULL, eyl.x, 0); unsigned cy = (unsigned)__shfl_sync(FULL, eyl.y, 0);
            unsigned Hp0 = 0, Hp1 = 0, Hp2 = 0, Hp3 = 0;
#pragma unroll 3
            for (int r = 0; r <= L3_RG; r++) {
                const unsigned X = __funnelshift_r(p0[r * L3_SP], p0[r * L3_SP + 1], sh), Y = __funnelshift_r(p0[r * L3_SP + 1], p0[r * L3_SP + 2], sh);
                const unsigned Hc0 = __dp2a_lo(c0, __byte_perm(X, Y, sel0), 0u) >> 4, Hc1 = __dp2a_lo(c1, __byte_perm(X, Y, sel1), 0u) >> 4;
                const unsigned Hc2 = __dp2a_lo(c2, __byte_perm(X, Y, sel2), 0u) >> 4, Hc3 = __dp2a_lo(c3, __byte_perm(X, Y, sel3), 0u) >> 4;
                if (r > 0 && sy == 8 * g8 + r - 1) {             // warp-uniform: destination row dy interpolates rows r-1 and r
                    const unsigned b0 = (cy & 0xffffu) << 16, b1 = (cy >> 16) << 16;
                    const unsigned o = ((__umulhi(b0, Hp0) + __umulhi(b1, Hc0) + 2u) >> 2) + (((__umulhi(b0, Hp1) + __umulhi(b1, Hc1) + 2u) >> 2) << 8) +
                                       (((__umulhi(b0, Hp2) + __umulhi(b1, Hc2) + 2u) >> 2) << 16) + (((__umulhi(b0, Hp3) + __umulhi(b1, Hc3) + 2u) >> 2) << 24);
                    if (active) *reinterpret_cast<uint32_t*>(dp) = o;
                    dp += a.dst_pitch;
                    dy++;
                    sy = __shfl_sync(FULL, eyl.x, (dy - dy0) & 31); cy = (unsigned)__shfl_sync(FULL, eyl.y, (dy - dy0) & 31);
                    if (dy >= dy1) sy = 0x7fffffff;
                }
                Hp0 = Hc0; Hp1 = Hc1; Hp2 = Hc2; Hp3 = Hc3;
            }
        }
    }

    // ---- 4. FAST quick reject on words -> warp-compacted candidate queue; score rows -1 .. TH of the tile ----
    const bool fast_tile = a.fast_on && (x0 + L3_TWPX >= 19) && (x0 - 1 <= a.w - 20) && (y0 + TH >= 19) && (y0 - 1 <= a.h - 20);
    if (fast_tile && role_fast) {
        // Conservative threshold t' = 2^k - 1 <= minTh (exact for the usual minTh = 7): |d| > t'  <=>  (|d| & ~t') != 0, so the
        // OR of two ring differences is tested with one mask; nz(x) = bit 7 of (x >> 1) + (0x80 - 2^(k-1)) for x with k low zero bits.
        const unsigned M = a.qmask, C = a.qadd;
        // valid bytes of this lane's word: pixels max(19, x0-1) .. min(w-20, x0+120)
        const int lo = max(19, x0 - 1) - (x0 - 4) - 4 * lw, hi = min(a.w - 20, x0 + L3_TWPX) - (x0 - 4) - 4 * lw;
        unsigned valid = 0x80808080u;
        if (lo > 3 || hi < 0) valid = 0;
        else {
            if (lo > 0) valid &= 0xffffffffu << (8 * lo);
            if (hi < 3) valid &= 0xffffffffu >> (8 * (3 - hi));
        }
        // Each warp scans 9 CONSECUTIVE score-map rows s = 9 * warp - 1 + it (the last warp has fewer), so the staged rows slide
        // through registers: a row is loaded once (own word + both neighbours) and its +-2-pixel and +-3-pixel shifted views are
        // formed once, whichever of the five roles (centre, +-2, +-3) it plays for the rows around it.
        constexpr int RPW9 = 9;
        static_assert(RPW9 * NW >= TH + 2, "9 rows per warp cover the tile and its ring");
        const int sb = RPW9 * warp - 1;                           // first score-map row of this warp (tile row index)
        // bit it: row s = sb + it is a scored row of the image (s <= TH, 19 <= y0 + s <= h - 20): a run of bits lo_it .. hi_it
        const int lo_it = max(0, 19 - y0 - sb), hi_it = min(min(RPW9 - 1, TH - sb), a.h - 20 - y0 - sb);
        const unsigned rowok = lo_it <= hi_it ? ((2u << hi_it) - 1u) & ~((1u << lo_it) - 1u) : 0u;
        // private candidate flags: byte b of m0 holds rows it = 0..7 of pixel b (bit it), m1 holds row 8 at bit 7 of byte b
        unsigned m0 = 0, m1 = 0;
        if (rowok) {                                              // warp-uniform
            const uint32_t* rp = tw + (sb + 1) * L3_SP + lw;     // staged row j of the window = tile row sb - 3 + j
            unsigned Vw[7], Aw[7] = {}, Bw[7] = {};              // sliding window: own word, +2-pixel view, -2-pixel view
#pragma unroll
            for (int j = 0; j < 6; j++) {
                const unsigned V = rp[j * L3_SP];
                Vw[j] = V;
                if (j >= 1) { Aw[j] = __funnelshift_r(V, rp[j * L3_SP + 1], 16); Bw[j] = __funnelshift_r(rp[j * L3_SP - 1], V, 16); }
            }
#pragma unroll
            for (int it = 0; it < RPW9; it++) {
                // window slots: 0 = row -3, 1 = row -2, 3 = centre, 5 = row +2, 6 = row +3 (entered now)
                Vw[6] = rp[(it + 6) * L3_SP];
                const unsigned V5 = Vw[5];
                if (it > 0) { Aw[5] = __funnelshift_r(V5, rp[(it + 5) * L3_SP + 1], 16); Bw[5] = __funnelshift_r(rp[(it + 5) * L3_SP - 1], V5, 16); }
                const unsigned V = Vw[3];
                const unsigned Cw = __funnelshift_r(V, rp[(it + 3) * L3_SP + 1], 24), Dw = __funnelshift_r(rp[(it + 3) * L3_SP - 1], V, 8);
                // any 9-arc contains two ADJACENT ones of the axis four  0 (0,+3), 4 (+3,0), 8 (0,-3), 12 (-3,0)
                // and two adjacent ones of the diagonal four             2 (+2,+2), 6 (+2,-2), 10 (-2,-2), 14 (-2,+2)
                const unsigned p08 = (__vabsdiffu4(V, Vw[6]) | __vabsdiffu4(V, Vw[0])) & M;
                const unsigned p4c = (__vabsdiffu4(V, Cw) | __vabsdiffu4(V, Dw)) & M;
                const unsigned p2a = (__vabsdiffu4(V, Aw[5]) | __vabsdiffu4(V, Bw[1])) & M;
                const unsigned p6e = (__vabsdiffu4(V, Aw[1]) | __vabsdiffu4(V, Bw[5])) & M;
                // minTh == 0: the host passes M = 0, C = 0x80 per byte, so every valid pixel is a candidate
                const unsigned cand = ((p08 >> 1) + C) & ((p4c >> 1) + C) & ((p2a >> 1) + C) & ((p6e >> 1) + C) & valid;
                if (it < 8) m0 = (m0 >> 1) + cand;               // after row 7: bit it of byte b = row it (no carry: bit 7 of m0 >> 1 is byte b + 1's
                else m1 = cand;                                   // bit 0, which only fills on the eighth shift)
#pragma unroll
                for (int j = 0; j < 6; j++) { Vw[j] = Vw[j + 1]; Aw[j] = Aw[j + 1]; Bw[j] = Bw[j + 1]; }
            }
            m0 &= (rowok & 0xffu) * 0x01010101u;                  // rows outside the scored range (image border tiles)
            if (!(rowok & 0x100u)) m1 = 0;
        }
        // each lane appends its own candidates: one shared atomic per lane, no warp scan (queue order is irrelevant)
        const int n = __popc(m0) + __popc(m1);
        if (n) {
            // queue code: row group << 12 | staged word << 6 | tt, tt = 8 * byte + it for the rows it = 0..7, 32 + byte for row 8 (decoded,
            // 32 candidates per instruction, by the scoring pass; this loop runs once per candidate of the warp's busiest lane)
            int slot = atomicAdd(&s_qn, n);
            const unsigned lanecode = ((unsigned)warp << 12) | ((unsigned)lw << 6);
            while (m0) {
                const int t = __ffs((int)m0) - 1;
                m0 &= m0 - 1;
                queue[slot++] = (uint16_t)(lanecode | (unsigned)t);
            }
            while (m1) {                                          // bit t = 8 * byte + 7
                const int t = __ffs((int)m1) - 1;
                m1 &= m1 - 1;
                queue[slot++] = (uint16_t)(lanecode | 32u | (unsigned)(t >> 3));
            }
        }
        if (OW == 32 && lane < 2 * RPW9) {
            // the ring pixels -1 and 128 of the warp's 9 rows lie in the halo words: one lane each, the same conservative test on single bytes
            const int it = lane >> 1, side = lane & 1;
            const int c = side ? 4 + L3_TWPX : 3;                 // staged byte column
            const int gxr = x0 - 4 + c;
            if (((rowok >> it) & 1u) && gxr >= 19 && gxr <= a.w - 20) {
                const uint8_t* cp = tb + (sb + it + 4) * L3_SPB + c;          // tile row sb + it is staged row sb + it + 4
                const int tq = (int)((~M) & 0xffu);               // t' (M = 0: minTh = 0, everything passes)
                const int v = cp[0];
                auto far = [&](int off) { return (int)(abs(v - (int)cp[off]) > tq); };         // branch-free: all eight loads go out together
                const int cand = (far(3 * L3_SPB) | far(-3 * L3_SPB)) & (far(3) | far(-3)) &
                                 (far(2 * L3_SPB + 2) | far(-2 * L3_SPB - 2)) & (far(-2 * L3_SPB + 2) | far(2 * L3_SPB - 2));
                if (cand || M == 0u)                              // staged byte column c = 4 * word + byte: word 0 byte 3, or word OW + 1 byte 0
                    queue[atomicAdd(&s_qn, 1)] = (uint16_t)(((unsigned)warp << 12) | ((unsigned)(c >> 2) << 6) | (it < 8 ? (unsigned)(8 * (c & 3) + it) : 32u + (c & 3)));
            }
        }
    }
    __syncthreads();
    if (!fast_tile) return;

    // ---- 5. exact arc score of the queued pixels.  Each warp owns a contiguous chunk of the queue and compacts the entries that
    //         survive (owned pixel, score above minTh) to the front of its own chunk: warp-level sync only. ----
    const int qn = s_qn;
    const int chunk = (qn + NWT - 1) / NWT;
    const int cbeg = wraw * chunk, cend = min(qn, cbeg + chunk);
    int wcnt = 0;
    for (int j0 = cbeg; j0 < cend; j0 += 32) {
        const int j = j0 + lane;
        bool surv = false; int pos = 0;
        if (j < cend) {
            const int code = queue[j], tt = code & 63;
            const int r = 9 * (code >> 12) + (tt < 32 ? (tt & 7) : 8);                   // score-map row (tile row + 1): row group base 9 * warp - 1, + it, + 1
            const int c = 4 * ((code >> 6) & 63) + (tt < 32 ? (tt >> 3) : tt - 32);     // staged byte column
            pos = (r << SCSH) | c;
            const int m = l3_arc_score<L3_SPB>(tb + (r + 3) * L3_SPB + c);
            if (m > a.minTh) {
                sc[r * L3_SCP + c] = (uint8_t)m;
                surv = r >= 1 && r <= TH && c >= 4 && c < 4 + L3_TWPX;     // the ring around the tile is scored for its neighbours only
            }
        }
        const unsigned bal = __ballot_sync(FULL, surv);          // (also orders this pass's queue reads before the writes below)
        if (surv) queue[cbeg + wcnt + __popc(bal & ((1u << lane) - 1))] = (uint16_t)pos;
        wcnt += __popc(bal);
    }
    __syncthreads();                                             // every score of the tile is in the map

    // ---- 6. cell-local 3x3 non-max suppression (strict '>') on the warp's survivors; kept corners go straight to the frame's
    //         record list (one global atomic per warp pass) ----
    uint32_t* __restrict__ rec = a.rec + (long long)f * a.rec_fstride;
    int* rec_count = a.rec_count + (long long)f * a.rec_count_stride;
    for (int j0 = cbeg; j0 < cbeg + wcnt; j0 += 32) {
        const int j = j0 + lane;
        bool keep = false; unsigned record = 0;
        if (j < cbeg + wcnt) {
            const int pos = queue[j];
            const int r = pos >> SCSH, c = pos & ((1 << SCSH) - 1);
            const uint8_t* sp = sc + r * L3_SCP + c;
            const int m = sp[0];
            const int gx = x0 - 4 + c, gy = y0 - 1 + r;
            // position inside the 30-px FAST cell: a neighbour in ANOTHER cell never suppresses (cv::FAST ran per cell)
            const int cellx = __umulhi((unsigned)(gx - 19), a.magicW), celly = __umulhi((unsigned)(gy - 19), a.magicH);
            const int rx = gx - 19 - cellx * a.wCell, ry = gy - 19 - celly * a.hCell;
            const bool left = rx > 0, right = rx < a.wCell - 1, up = ry > 0, down = ry < a.hCell - 1;
            keep = true;
            if (up) { if ((left && sp[-L3_SCP - 1] >= m) || sp[-L3_SCP] >= m || (right && sp[-L3_SCP + 1] >= m)) keep = false; }
            if ((left && sp[-1] >= m) || (right && sp[1] >= m)) keep = false;
            if (down) { if ((left && sp[L3_SCP - 1] >= m) || sp[L3_SCP] >= m || (right && sp[L3_SCP + 1] >= m)) keep = false; }
            record = ((unsigned)gy << 20) | ((unsigned)gx << 8) | (unsigned)m;
        }
        const unsigned bal = __ballot_sync(FULL, keep);
        if (bal) {
            int base = 0;
            if (lane == 0) base = atomicAdd(rec_count, __popc(bal));
            base = __shfl_sync(FULL, base, 0) + __popc(bal & ((1u << lane) - 1));
            if (keep && base < a.rec_cap) rec[base] = record;
        }
    }
}

// Generic bilinear resize (one thread per destination pixel, reads global memory): used instead of the fused path when
// the scale factor is too large for the 4-pixel halo (scaleFactor > 1.25).  Same arithmetic as step 3 above.
__global__ void resize_plain_kernel(const uint8_t* __restrict__ src, long long src_pitch, long long src_fstride, int sw, int sh,
                                    uint8_t* __restrict__ dst, long long dst_pitch, long long dst_fstride, int dw, int dh,
                                    const int2* __restrict__ xt, const int2* __restrict__ yt)
{
    const int dx = blockIdx.x * blockDim.x + threadIdx.x, dy = blockIdx.y;
    if (dx >= dw || dy >= dh) return;
    const uint8_t* s = src + (long long)blockIdx.z * src_fstride;
    const int2 ex = xt[dx], ey = yt[dy];
    const int x1 = min(ex.x + 1, sw - 1), y1 = min(ey.x + 1, sh - 1);
    const unsigned ca = (unsigned)ex.y & 0xffffu, cb = (unsigned)ex.y >> 16;
    const unsigned b0 = ((unsigned)ey.y & 0xffffu) << 16, b1 = ((unsigned)ey.y >> 16) << 16;
    const uint8_t* r0 = s + (long long)ey.x * src_pitch; const uint8_t* r1 = s + (long long)y1 * src_pitch;
    const unsigned H0 = (r0[ex.x] * ca + r0[x1] * cb) >> 4, H1 = (r1[ex.x] * ca + r1[x1] * cb) >> 4;
    dst[(long long)blockIdx.z * dst_fstride + (long long)dy * dst_pitch + dx] = (uint8_t)((__umulhi(b0, H0) + __umulhi(b1, H1) + 2u) >> 2);
}

}  // namespace orbb
