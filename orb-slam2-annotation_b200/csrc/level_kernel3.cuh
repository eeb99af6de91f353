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

namespace orbb {

constexpr int L3_LANES = 32;                 // staged word columns: halo word | 30 owned words | halo word
constexpr int L3_OWNW = 30;                  // owned words per tile row
constexpr int L3_TWPX = 4 * L3_OWNW;         // 120 owned pixels per tile row
constexpr int L3_RG = 8;                     // rows per warp
constexpr int L3_SP = 34;                    // shared tile pitch in words (1 pad word on each side of the 32 staged)
constexpr int L3_SPB = 4 * L3_SP;            // ... in bytes
constexpr int L3_SCP = 128;                  // score map pitch (byte column = staged tile byte column)

struct Level3Args {
    const uint8_t* src; long long src_pitch, src_fstride;   // level L (external input image when L == 0)
    int w, h;
    uint8_t* blur; long long blur_pitch, blur_fstride;      // blurred level L out
    uint8_t* dst; long long dst_pitch, dst_fstride;         // level L+1 out (null: last level, or resize done elsewhere)
    const int2* xt;          // per destination x (padded to a multiple of 4): {source x, a0 | a1 << 16}
    const int2* yt;          // per destination y: {source y, b0 | b1 << 16}
    const int* xws;          // [tiles_x + 1] first destination WORD owned by each tile column
    const int* yrs;          // [ceil(h / 8) + 1] first destination row whose source row lies in each 8-row group
    int fast_on, wCell, hCell, minTh;
    unsigned magicW, magicH;                                // ceil(2^32 / wCell), ceil(2^32 / hCell)
    uint32_t* rec; int rec_cap; long long rec_fstride;      // raw FAST records  y<<20 | x<<8 | m
    int* rec_count; int rec_count_stride;
    int q[7];                                               // Q8 blur kernel
    int legacy;                                             // 1: kernel sums to 257 -> the result needs a clamp
};

__device__ __forceinline__ int l3_reflect101(int p, int n)
{
    if (p < 0) p = -p;
    if (p >= n) p = 2 * (n - 1) - p;
    return min(max(p, 0), n - 1);
}

// per byte: bit 7 set where x > t (0 <= t <= 126); other bits are garbage (callers mask with 0x80808080)
__device__ __forceinline__ unsigned l3_gt_raw(unsigned x, unsigned t1 /* (t+1) * 0x01010101 */)
{
    return (((x & 0x7f7f7f7fu) | 0x80808080u) - t1) | x;
}

// Threshold-independent FAST arc score m = max over the 16 arcs of 9 contiguous ring pixels of
// max(min(v - p_k), min(p_k - v)); cv::FAST's response is m - 1 and a pixel is a corner at threshold t iff m > t.
// Both polarities ride in one register as biased u16 halves: lo = 256 + v - p_k, hi = 256 - v + p_k.
__device__ __forceinline__ int l3_arc_score(const uint8_t* c)
{
    constexpr int P = L3_SPB;
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
    unsigned best = 0;
#pragma unroll
    for (int k = 0; k < 16; k += 2) {
        const unsigned a9 = __vimin3_u16x2(a3[k], a3[(k + 3) & 15], a3[(k + 6) & 15]);
        const unsigned b9 = __vimin3_u16x2(a3[k + 1], a3[(k + 4) & 15], a3[(k + 7) & 15]);
        best = __vimax3_u16x2(best, a9, b9);
    }
    return (int)max(best & 0xffffu, best >> 16) - 256;
}

__device__ __forceinline__ unsigned l3_pack4(int b0, int b1, int b2, int b3)
{
    return (unsigned)b0 | ((unsigned)b1 << 8) | ((unsigned)b2 << 16) | ((unsigned)b3 << 24);
}

template <int NW>
__global__ void __launch_bounds__(32 * NW) level3_kernel(const __grid_constant__ Level3Args a)
{
    constexpr int TH = L3_RG * NW, SMH = TH + 8, SCH = TH + 2, NT = 32 * NW;
    constexpr int QCAP = SCH * 122;                      // every scored pixel a candidate (uniform-noise images)
    constexpr int EMIT_CAP = (TH / 2 + 3) * 65;          // NMS survivors are never 8-adjacent inside a cell
    constexpr unsigned FULL = 0xffffffffu;
    __shared__ __align__(16) uint32_t tw[SMH * L3_SP];   // tile + halo, 4 pixels per word, word (r, c) at r*SP + 1 + c
    __shared__ __align__(4) uint8_t sc[SCH * L3_SCP];    // arc scores (> minTh) of the tile's pixels + 1-px ring
    __shared__ uint16_t queue[QCAP];
    __shared__ uint32_t emit[EMIT_CAP];
    __shared__ int s_qn, s_en, s_ebase;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int f = blockIdx.z;
    const int x0 = blockIdx.x * L3_TWPX, y0 = blockIdx.y * TH;
    const uint8_t* __restrict__ src = a.src + (long long)f * a.src_fstride;
    const uint8_t* tb = reinterpret_cast<const uint8_t*>(tw);
    if (tid == 0) { s_qn = 0; s_en = 0; }

    // ---- 1. stage tile + halo: lane = word column; REFLECT_101 at the image border (what the blur of a cloned level sees) ----
    {
        const bool word_ok = ((a.src_pitch & 3) == 0) && ((reinterpret_cast<unsigned long long>(src) & 3) == 0);
        const int gx = x0 - 4 + 4 * lane;
        const bool inside = word_ok && gx >= 0 && gx + 3 < a.w;
#pragma unroll
        for (int r = warp; r < SMH; r += NW) {
            const int gy = l3_reflect101(y0 - 4 + r, a.h);
            const uint8_t* row = src + (long long)gy * a.src_pitch;
            unsigned v = 0;
            if (inside) v = *reinterpret_cast<const uint32_t*>(row + gx);
            else if (gx < a.w + 4) {
#pragma unroll
                for (int b = 0; b < 4; b++) v |= (unsigned)row[l3_reflect101(gx + b, a.w)] << (8 * b);
            }
            tw[r * L3_SP + 1 + lane] = v;
            if (lane < 2) tw[r * L3_SP + lane * (L3_SP - 1)] = 0;       // pad words: read (and masked out) by the edge lanes
        }
    }
    for (int i = tid; i < SCH * L3_SCP / 4; i += NT) reinterpret_cast<uint32_t*>(sc)[i] = 0;
    __syncthreads();

    const int ry0 = L3_RG * warp;                 // this warp's rows inside the tile
    const bool rows_live = y0 + ry0 < a.h;

    // ---- 2. blur: vertical pass on u16x2 pairs (even pixels | odd pixels), horizontal pass with IDP.2A ----
    if (rows_live) {
        unsigned E[14], O[14];
#pragma unroll
        for (int i = 0; i < 14; i++) {
            const unsigned w = tw[(ry0 + 1 + i) * L3_SP + 1 + lane];
            E[i] = w & 0x00ff00ffu;
            O[i] = __byte_perm(w, 0u, 0x4341);
        }
        const unsigned q0 = a.q[0], q1 = a.q[1], q2 = a.q[2], q3 = a.q[3];      // symmetric kernel (checked by the host)
        const unsigned KA = l3_pack4(q0, q2, 0, q1), KB = l3_pack4(q3, q1, q2, q0), KC = l3_pack4(0, q0, 0, q1);
        const unsigned KD = l3_pack4(q2, q2, q3, q1), KE = l3_pack4(q0, 0, q1, 0), KF = l3_pack4(q1, q3, q2, q2);
        const unsigned KG = l3_pack4(q0, q2, q1, q3), KH = l3_pack4(q2, q0, q1, 0);
        uint8_t* __restrict__ bl = a.blur + (long long)f * a.blur_fstride;
        const int gx = x0 - 4 + 4 * lane;
        const bool store_lane = lane >= 1 && lane <= L3_OWNW && gx < a.w;
#pragma unroll
        for (int j = 0; j < L3_RG; j++) {
            const unsigned VE = q3 * E[j + 3] + q2 * (E[j + 2] + E[j + 4]) + q1 * (E[j + 1] + E[j + 5]) + q0 * (E[j] + E[j + 6]);
            const unsigned VO = q3 * O[j + 3] + q2 * (O[j + 2] + O[j + 4]) + q1 * (O[j + 1] + O[j + 5]) + q0 * (O[j] + O[j + 6]);
            const unsigned lE = __shfl_up_sync(FULL, VE, 1), lO = __shfl_up_sync(FULL, VO, 1);
            const unsigned rE = __shfl_down_sync(FULL, VE, 1), rO = __shfl_down_sync(FULL, VO, 1);
            // taps: q0..q6 with q4=q2, q5=q1, q6=q0; pairs are (even, odd) = (V0|V2<<16, V1|V3<<16) of the left / own / right word
            unsigned r0 = __dp2a_lo(lO, KA, 32768u); r0 = __dp2a_hi(lE, KA, r0); r0 = __dp2a_lo(VE, KB, r0); r0 = __dp2a_hi(VO, KB, r0);
            unsigned r1 = __dp2a_lo(lE, KC, 32768u); r1 = __dp2a_hi(lO, KC, r1); r1 = __dp2a_lo(VE, KD, r1); r1 = __dp2a_hi(VO, KD, r1);
            r1 = __dp2a_lo(rE, KE, r1);
            unsigned r2 = __dp2a_lo(lO, KC, 32768u); r2 = __dp2a_lo(VE, KF, r2); r2 = __dp2a_hi(VO, KF, r2); r2 = __dp2a_hi(rE, KE, r2);
            r2 = __dp2a_lo(rO, KE, r2);
            unsigned r3 = __dp2a_lo(VE, KG, 32768u); r3 = __dp2a_hi(VO, KG, r3); r3 = __dp2a_lo(rE, KH, r3); r3 = __dp2a_hi(rO, KH, r3);
            unsigned o;
            if (!a.legacy) {        // sums < 2^24: the rounded result is byte 2
                o = __byte_perm(__byte_perm(r0, r1, 0x0062), __byte_perm(r2, r3, 0x0062), 0x5410);
            } else {
                o = min(r0 >> 16, 255u) | (min(r1 >> 16, 255u) << 8) | (min(r2 >> 16, 255u) << 16) | (min(r3 >> 16, 255u) << 24);
            }
            const int gy = y0 + ry0 + j;
            if (store_lane && gy < a.h) *reinterpret_cast<uint32_t*>(bl + (long long)gy * a.blur_pitch + gx) = o;
        }
    }

    // ---- 3. bilinear resize: lane = one destination word whose first source column lies in this tile, rows of this warp ----
    if (a.dst && rows_live) {
        uint8_t* __restrict__ dst = a.dst + (long long)f * a.dst_fstride;
        const int k1 = a.xws[blockIdx.x + 1];
        const int g8 = (y0 + ry0) >> 3;
        const int dy0 = a.yrs[g8], dy1 = a.yrs[g8 + 1];
        for (int k = a.xws[blockIdx.x] + lane; k < k1; k += 32) {
            int off[4]; unsigned ca[4], cb[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int2 e = a.xt[4 * k + j];
                off[j] = e.x - (x0 - 4) + 4;                    // byte offset of the left source pixel inside a staged row
                ca[j] = (unsigned)e.y & 0xffffu; cb[j] = (unsigned)e.y >> 16;
            }
            int have = -100000;                                  // staged row whose interpolation sits in H1
            unsigned H0[4], H1[4] = {0, 0, 0, 0};
            for (int dy = dy0; dy < dy1; dy++) {
                const int2 ey = a.yt[dy];
                const int rr = ey.x - (y0 - 4);
                const unsigned b0 = ((unsigned)ey.y & 0xffffu) << 16, b1 = ((unsigned)ey.y >> 16) << 16;
                const uint8_t* p0 = tb + rr * L3_SPB;
                if (rr == have) {
#pragma unroll
                    for (int j = 0; j < 4; j++) H0[j] = H1[j];
                } else {
#pragma unroll
                    for (int j = 0; j < 4; j++) H0[j] = (p0[off[j]] * ca[j] + p0[off[j] + 1] * cb[j]) >> 4;
                }
#pragma unroll
                for (int j = 0; j < 4; j++) H1[j] = (p0[L3_SPB + off[j]] * ca[j] + p0[L3_SPB + off[j] + 1] * cb[j]) >> 4;
                have = rr + 1;
                unsigned o = 0;
#pragma unroll
                for (int j = 0; j < 4; j++) o |= ((__umulhi(b0, H0[j]) + __umulhi(b1, H1[j]) + 2u) >> 2) << (8 * j);
                *reinterpret_cast<uint32_t*>(dst + (long long)dy * a.dst_pitch + 4 * k) = o;
            }
        }
    }

    // ---- 4. FAST quick reject on words -> warp-compacted candidate queue; score rows -1 .. TH of the tile ----
    const bool fast_tile = a.fast_on && (x0 + L3_TWPX >= 19) && (x0 - 1 <= a.w - 20) && (y0 + TH >= 19) && (y0 - 1 <= a.h - 20);
    if (fast_tile) {
        const unsigned t1 = (unsigned)(min(a.minTh, 126) + 1) * 0x01010101u;
        // valid bytes of this lane's word: pixels max(19, x0-1) .. min(w-20, x0+120)
        const int lo = max(19, x0 - 1) - (x0 - 4) - 4 * lane, hi = min(a.w - 20, x0 + L3_TWPX) - (x0 - 4) - 4 * lane;
        unsigned valid = 0x80808080u;
        if (lo > 3 || hi < 0) valid = 0;
        else {
            if (lo > 0) valid &= 0xffffffffu << (8 * lo);
            if (hi < 3) valid &= 0xffffffffu >> (8 * (3 - hi));
        }
        for (int s = warp - 1; s <= TH; s += NW) {
            const int gy = y0 + s;
            if (gy < 19 || gy > a.h - 20) continue;                                  // warp-uniform
            const uint32_t* rowp = tw + (s + 4) * L3_SP + 1 + lane;
            const unsigned V = rowp[0], Lw = rowp[-1], Rw = rowp[1];
            // ring pixels 0 (0,+3), 8 (0,-3), 4 (+3,0), 12 (-3,0): any 9-arc contains two ADJACENT ones of these four
            const unsigned g0 = l3_gt_raw(__vabsdiffu4(V, rowp[3 * L3_SP]), t1), g8 = l3_gt_raw(__vabsdiffu4(V, rowp[-3 * L3_SP]), t1);
            const unsigned g4 = l3_gt_raw(__vabsdiffu4(V, __funnelshift_r(V, Rw, 24)), t1);
            const unsigned g12 = l3_gt_raw(__vabsdiffu4(V, __funnelshift_r(Lw, V, 8)), t1);
            unsigned cand = (g0 | g8) & (g4 | g12) & valid;
            if (cand) {
                // ... and two adjacent ones of the diagonal four: 2 (+2,+2), 6 (+2,-2), 10 (-2,-2), 14 (-2,+2)
                const unsigned Vd = rowp[2 * L3_SP], Ld = rowp[2 * L3_SP - 1], Rd = rowp[2 * L3_SP + 1];
                const unsigned Vu = rowp[-2 * L3_SP], Lu = rowp[-2 * L3_SP - 1], Ru = rowp[-2 * L3_SP + 1];
                const unsigned g2 = l3_gt_raw(__vabsdiffu4(V, __funnelshift_r(Vd, Rd, 16)), t1);
                const unsigned g14 = l3_gt_raw(__vabsdiffu4(V, __funnelshift_r(Ld, Vd, 16)), t1);
                const unsigned g6 = l3_gt_raw(__vabsdiffu4(V, __funnelshift_r(Vu, Ru, 16)), t1);
                const unsigned g10 = l3_gt_raw(__vabsdiffu4(V, __funnelshift_r(Lu, Vu, 16)), t1);
                cand &= (g2 | g10) & (g6 | g14);
            }
            if (__any_sync(FULL, cand != 0)) {
                const int n = (int)(((cand >> 7) * 0x01010101u) >> 24);
                int incl = n;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
                int base = 0;
                if (lane == 31) base = atomicAdd(&s_qn, incl);
                base = __shfl_sync(FULL, base, 31);
                int slot = base + incl - n;
                const int pos = (s + 1) * L3_SCP + 4 * lane;
#pragma unroll
                for (int b = 0; b < 4; b++)
                    if (cand & (0x80u << (8 * b))) queue[slot++] = (uint16_t)(pos + b);
            }
        }
    }
    __syncthreads();
    if (!fast_tile) return;

    // ---- 5. exact arc score of the queued pixels ----
    const int qn = s_qn;
    for (int j = tid; j < qn; j += NT) {
        const int pos = queue[j];
        const int r = pos >> 7, c = pos & 127;                    // score-map row (tile row + 1), staged byte column
        const int m = l3_arc_score(tb + (r + 3) * L3_SPB + 4 + c);
        sc[pos] = (uint8_t)(m > a.minTh ? m : 0);
    }
    __syncthreads();

    // ---- 6. cell-local 3x3 non-max suppression (strict '>') and emission ----
    for (int j = tid; j < qn; j += NT) {
        const int pos = queue[j];
        const int r = pos >> 7, c = pos & 127;
        if (r < 1 || r > TH || c < 4 || c >= 4 + L3_TWPX) continue;     // the ring around the tile is scored for its neighbours only
        const uint8_t* sp = sc + pos;
        const int m = sp[0];
        if (!m) continue;
        const int gx = x0 - 4 + c, gy = y0 - 1 + r;
        // position inside the 30-px FAST cell: a neighbour in ANOTHER cell never suppresses (cv::FAST ran per cell)
        const int cellx = __umulhi((unsigned)(gx - 19), a.magicW), celly = __umulhi((unsigned)(gy - 19), a.magicH);
        const int rx = gx - 19 - cellx * a.wCell, ry = gy - 19 - celly * a.hCell;
        const bool left = rx > 0, right = rx < a.wCell - 1, up = ry > 0, down = ry < a.hCell - 1;
        bool keep = true;
        if (up) { if ((left && sp[-L3_SCP - 1] >= m) || sp[-L3_SCP] >= m || (right && sp[-L3_SCP + 1] >= m)) keep = false; }
        if ((left && sp[-1] >= m) || (right && sp[1] >= m)) keep = false;
        if (down) { if ((left && sp[L3_SCP - 1] >= m) || sp[L3_SCP] >= m || (right && sp[L3_SCP + 1] >= m)) keep = false; }
        if (keep) {
            const int e = atomicAdd(&s_en, 1);
            if (e < EMIT_CAP) emit[e] = ((unsigned)gy << 20) | ((unsigned)gx << 8) | (unsigned)m;
        }
    }
    __syncthreads();
    const int en = min(s_en, EMIT_CAP);
    if (en == 0) return;
    if (tid == 0) s_ebase = atomicAdd(a.rec_count + (long long)f * a.rec_count_stride, en);
    __syncthreads();
    const int ebase = s_ebase;
    uint32_t* __restrict__ rec = a.rec + (long long)f * a.rec_fstride;
    for (int j = tid; j < en; j += NT)
        if (ebase + j < a.rec_cap) rec[ebase + j] = emit[j];
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
